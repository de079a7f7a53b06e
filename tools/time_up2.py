"""Times the 2x f32 upscale (K5 shape) with the register-window kernel and the tile kernels.   python tools/time_up2.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from image_b200 import FilterType  # noqa: E402
from image_b200 import _lib as L  # noqa: E402

lib = L.lib()
st = torch.cuda.current_stream().cuda_stream
hbm = 6551.0


def t(fn, n=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


for (w, h, FILT) in [(8192, 8192, FilterType.Lanczos3), (8192, 8192, FilterType.CatmullRom), (8192, 8192, FilterType.Triangle), (4096, 4096, FilterType.CatmullRom)]:
    src = torch.rand((1, h, w, 4), device="cuda")
    dst = torch.empty((1, 2 * h, 2 * w, 4), device="cuda")
    alg = (w * h + 4 * w * h) * 16
    for name, flags in [("up2", L.FP_FAST), ("general", L.FP_FAST | L.RESIZE_GENERAL)]:
        row = []
        envs = [None] if name == "general" else [None] + [("B200IMG_U2_BANDS", str(b)) for b in (30, 60, 120)]
        for env in envs:
            os.environ.pop("B200IMG_U2_BANDS", None)
            os.environ.pop("B200IMG_U2_TWS", None)
            if env:
                os.environ[env[0]] = env[1]
            ms = t(lambda: L.check(lib.b200img_resize_batch_dev(src.data_ptr(), 0, w, h, 4, L.F32, dst.data_ptr(), 0, 2 * w, 2 * h, 1,
                                                                int(FILT), 1.0, flags, st)))
            row.append(f"{env[0][-5:] + '=' + env[1] if env else 'default'} {ms:.4f} ms {alg / ms / 1e6:7.1f} GB/s {alg / ms / 1e6 / hbm:.3f}")
        print(f"{w}x{h} {FILT.name} {name}: " + " | ".join(row), flush=True)
    del src, dst
