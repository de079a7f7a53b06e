"""A small run of every kernel family for compute-sanitizer (memcheck / racecheck / initcheck / synccheck):
python tools/sanitize_small.py.  Small shapes, all sample types, the fused kernels forced; results are checked
against the oracle too, so a sanitizer-clean run is also a correct one."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import oracle as o  # noqa: E402
from image_b200 import FilterType, GaussianBlurParameters, imageops  # noqa: E402
from image_b200 import _lib as L  # noqa: E402
from image_b200.multi import run_bands  # noqa: E402

dev = torch.device("cuda:0")
rng = np.random.default_rng(1)
for dt in (np.uint8, np.uint16, np.float32):
    img = rng.random((70, 210, 3), dtype=np.float32) if dt == np.float32 else rng.integers(0, np.iinfo(dt).max + 1, (70, 210, 3), dtype=dt)
    d = torch.from_numpy(img).to(dev)
    for sigma in (1.2, 10.0):
        p = GaussianBlurParameters.new_from_sigma(sigma)
        for fl in (L.GAUSS_FUSED, L.GAUSS_TWO_KERNELS):
            got = imageops.blur_advanced(d, p, flags=fl).cpu().numpy()
            want = o.gaussian_blur(img, o.gauss_params_from_sigma(sigma), "generic")
            assert np.allclose(got.astype(np.float64), want.astype(np.float64), rtol=1e-5, atol=1.0 if dt != np.float32 else 1e-6)
        if dt != np.float32:
            assert np.array_equal(imageops.blur_dyn(d, p, flags=L.GAUSS_FUSED).cpu().numpy(), o.gaussian_blur(img, o.gauss_params_from_sigma(sigma), "dyn"))
    got = imageops.fast_blur(d, 3.0).cpu().numpy()
    want = o.fast_blur(img, 3.0)
    assert np.array_equal(got, want) if dt != np.float32 else np.allclose(got, want, rtol=1e-5, atol=2e-6)
    for f, name, (nw, nh) in ((FilterType.Lanczos3, "lanczos3", (105, 35)), (FilterType.CatmullRom, "catmullrom", (300, 100)), (FilterType.Nearest, "nearest", (50, 90))):
        assert np.array_equal(imageops.resize(d, nw, nh, f, flags=L.FP_STRICT).cpu().numpy(), o.resize(img, nw, nh, name))
        imageops.resize(d, nw, nh, f).cpu()
    assert np.array_equal(imageops.convert_pixels(d, 4, np.uint16).cpu().numpy(), o.convert(img.reshape(-1, 3), 3, np.uint16, 4).reshape(70, 210, 4))
    assert np.array_equal(imageops.thumbnail(d, 31, 19).cpu().numpy(), o.thumbnail(img, 31, 19)) or dt == np.float32
img = rng.integers(0, 256, (90, 64, 4), dtype=np.uint8)
assert np.array_equal(run_bands("gaussian", img, [0, 0, 0], params=GaussianBlurParameters.new_from_sigma(2.0), path=L.BLUR_DYN),
                      o.gaussian_blur(img, o.gauss_params_from_sigma(2.0), "dyn"))
assert np.array_equal(run_bands("fast_blur", img, [0, 0], sigma=2.0), o.fast_blur(img, 2.0))
assert np.array_equal(run_bands("resize", img, [0, 0, 0], nw=100, nh=50, filter=FilterType.Triangle, flags=L.FP_STRICT), o.resize(img, 100, 50, "triangle"))
torch.cuda.synchronize()
print("sanitize_small ok")
