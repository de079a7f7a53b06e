"""One-off stress of the fixed-window resamplers: many random shapes (batches, so that the kernels are picked) against the oracle.
python tools/stress_fixed_window.py [cases] [seed]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import oracle as o  # noqa: E402
from image_b200 import FilterType  # noqa: E402
from image_b200 import _lib as L  # noqa: E402

os.environ["B200IMG_U2_FORCE"] = "1"
lib = L.lib()
st = torch.cuda.current_stream().cuda_stream
cases = int(sys.argv[1]) if len(sys.argv) > 1 else 120
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 1)
filters = [FilterType.Triangle, FilterType.CatmullRom, FilterType.Gaussian, FilterType.Lanczos3]
DT = {np.uint8: L.U8, np.uint16: L.U16, np.float32: L.F32}
bad = 0
for it in range(cases):
    kind = rng.choice(["down", "down", "up"])
    f = filters[int(rng.integers(0, 4))]
    dtype = [np.uint8, np.uint8, np.uint16, np.float32][int(rng.integers(0, 4))]
    acc = float(rng.choice([0.0, 1.0]))
    if kind == "up":
        c, w, h, n = 4, int(rng.integers(16, 700)), int(rng.integers(16, 500)), int(rng.integers(1, 6))
        nw, nh = 2 * w, 2 * h
    else:
        r = int(rng.choice([2, 3, 4]))
        c = 4 if (r == 3 or dtype != np.uint8) else int(rng.integers(1, 5))
        nw, nh, n = int(rng.integers(8, 700)), int(rng.integers(8, 400)), int(rng.integers(1, 24))
        while (nw * r * c) % 4:
            nw += 1
        w, h = nw * r, nh * r
    if dtype == np.float32:
        imgs = rng.random((n, h, w, c), dtype=np.float32)
    else:
        imgs = rng.integers(0, np.iinfo(dtype).max + 1, (n, h, w, c), dtype=dtype)
    src = torch.from_numpy(imgs).cuda()
    dst = torch.empty((n, nh, nw, c), dtype=src.dtype, device="cuda")
    L.check(lib.b200img_resize_batch_dev(src.data_ptr(), 0, w, h, c, DT[dtype], dst.data_ptr(), 0, nw, nh, n, int(f), acc, L.FP_FAST, st))
    got = dst.cpu().numpy()
    for i in {0, n - 1}:
        want = o.resize(imgs[i], nw, nh, int(f), acc_init=acc)
        if dtype == np.float32:
            ok = np.allclose(got[i], want, rtol=1e-5, atol=1e-6)
        else:
            ok = np.abs(got[i].astype(np.int64) - want.astype(np.int64)).max() <= 1
        if not ok:
            bad += 1
            print("MISMATCH", kind, f.name, dtype.__name__, (n, w, h, c), "->", (nw, nh), "acc", acc, "image", i, flush=True)
print(f"{cases} cases, {bad} mismatches")
