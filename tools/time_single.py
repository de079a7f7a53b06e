"""Single-image latencies of the resampler: fixed-window kernels against the table-driven ones.   python tools/time_single.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from image_b200 import FilterType  # noqa: E402
from image_b200 import _lib as L  # noqa: E402

lib = L.lib()
st = torch.cuda.current_stream().cuda_stream


def t(fn, n=50):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3


CASES = [(1, 1920, 1080, 4, torch.uint8, 960, 540, FilterType.Lanczos3), (1, 1920, 1080, 3, torch.uint8, 960, 540, FilterType.Lanczos3),
         (1, 3840, 2160, 4, torch.uint8, 1920, 1080, FilterType.Lanczos3), (1, 1024, 1024, 4, torch.uint8, 256, 256, FilterType.CatmullRom),
         (2, 1920, 1080, 4, torch.uint8, 960, 540, FilterType.Lanczos3), (4, 1920, 1080, 4, torch.uint8, 960, 540, FilterType.Lanczos3),
         (1, 1920, 1080, 4, torch.uint8, 3840, 2160, FilterType.CatmullRom), (16, 1920, 1080, 4, torch.uint8, 3840, 2160, FilterType.CatmullRom),
         (4, 1920, 1080, 4, torch.uint16, 3840, 2160, FilterType.Lanczos3), (1, 1920, 1080, 4, torch.float32, 3840, 2160, FilterType.Lanczos3), (1, 960, 540, 4, torch.float32, 1920, 1080, FilterType.CatmullRom)]
for n, w, h, c, dt, nw, nh, f in CASES:
    src = (torch.rand((n, h, w, c), device="cuda") * (255 if dt == torch.uint8 else 65535 if dt == torch.uint16 else 1)).to(dt)
    dst = torch.empty((n, nh, nw, c), dtype=dt, device="cuda")
    ldt = L.U8 if dt == torch.uint8 else L.U16 if dt == torch.uint16 else L.F32
    row = []
    for name, flags, env in [("default", L.FP_FAST, None), ("general", L.FP_FAST | L.RESIZE_GENERAL, None)] + \
                            [("forced", L.FP_FAST, -1)]:
        for k in ("B200IMG_RI_BAND", "B200IMG_U2_BANDS"):
            os.environ.pop(k, None)
        os.environ.pop("B200IMG_U2_FORCE", None)
        if env:
            os.environ["B200IMG_U2_FORCE"] = "1"
        us = t(lambda: L.check(lib.b200img_resize_batch_dev(src.data_ptr(), 0, w, h, c, ldt, dst.data_ptr(), 0, nw, nh, n, int(f), 1.0, flags, st)))
        row.append(f"{name} {us:.1f}")
    print(f"{n}x{w}x{h}x{c} {str(dt)[6:]} -> {nw}x{nh} {f.name}: " + " | ".join(row) + " us", flush=True)
