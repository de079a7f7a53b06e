"""One-off stress of the fused box blur (bit-exact) and the fused Gaussian (+-1 LSB) on random shapes against the oracle.
python tools/stress_blur.py [cases] [seed]"""
import ctypes as C
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import oracle as o  # noqa: E402
from image_b200 import GaussianBlurParameters  # noqa: E402
from image_b200 import _lib as L  # noqa: E402

lib = L.lib()
st = torch.cuda.current_stream().cuda_stream
cases = int(sys.argv[1]) if len(sys.argv) > 1 else 150
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 2)
bad = 0
for it in range(cases):
    c = int(rng.integers(1, 5))
    w, h, n = int(rng.integers(8, 900)), int(rng.integers(8, 500)), int(rng.integers(1, 5))
    while (w * c) % 4:
        w += 1
    sigma = float(rng.choice([0.7, 1.5, 3.0, 6.0, 10.0, 17.0]))
    imgs = rng.integers(0, 256, (n, h, w, c), dtype=np.uint8)
    src = torch.from_numpy(imgs).cuda()
    dst = torch.empty_like(src)
    L.check(lib.b200img_fast_blur_batch_dev(src.data_ptr(), w, h, c, L.U8, dst.data_ptr(), n, sigma, st))
    got = dst.cpu().numpy()
    for i in {0, n - 1}:
        if not np.array_equal(got[i], o.fast_blur(imgs[i], sigma)):
            bad += 1
            print("MISMATCH fast_blur", (n, w, h, c), sigma, i, flush=True)
    if it % 3 == 0:
        p = GaussianBlurParameters.new_from_sigma(sigma)
        for flags in (L.GAUSS_FUSED, 0):
            L.check(lib.b200img_gaussian_blur_batch_dev(src.data_ptr(), w, h, c, L.U8, dst.data_ptr(), n, C.byref(p.raw), L.BLUR_GENERIC, flags, st))
            got = dst.cpu().numpy()
            want = o.gaussian_blur(imgs[n - 1], o.gauss_params_from_sigma(sigma), "generic", chunked=True)
            if np.abs(got[n - 1].astype(np.int64) - want.astype(np.int64)).max() > 1:
                bad += 1
                print("MISMATCH gaussian", (n, w, h, c), sigma, flags, flush=True)
print(f"{cases} cases, {bad} mismatches")
