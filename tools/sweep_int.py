"""Sweeps the strip width / band height of the fixed-window resampler (env B200IMG_RI_TW / B200IMG_RI_BAND) on K1 and K4."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from image_b200 import FilterType  # noqa: E402
from image_b200 import _lib as L  # noqa: E402

lib = L.lib()
st = torch.cuda.current_stream().cuda_stream


def t(fn, n=20):
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


CASES = [("K1", 64, 1920, 1080, 4, 960, 540, FilterType.Lanczos3, [56, 64, 72, 80, 88, 96, 104], [60, 96, 120, 180]),
         ("K4", 256, 1024, 1024, 4, 256, 256, FilterType.CatmullRom, [20, 28, 32, 36, 44, 52], [64, 112, 128, 256]),
         ("rgb8", 64, 1920, 1080, 3, 960, 540, FilterType.Lanczos3, [64, 80, 96, 112, 128, 160], [60, 120])]
for name, n, w, h, c, nw, nh, f, tws, bands in CASES:
    src = torch.randint(0, 256, (n, h, w, c), dtype=torch.int32, device="cuda").to(torch.uint8)
    dst = torch.empty((n, nh, nw, c), dtype=torch.uint8, device="cuda")
    call = lambda: L.check(lib.b200img_resize_batch_dev(src.data_ptr(), 0, w, h, c, L.U8, dst.data_ptr(), 0, nw, nh, n, int(f), 1.0, L.FP_FAST, st))
    os.environ.pop("B200IMG_RI_TW", None)
    os.environ.pop("B200IMG_RI_BAND", None)
    print(f"{name} default {t(call):.4f} ms", flush=True)
    for tw in tws:
        row = []
        for b in bands:
            os.environ["B200IMG_RI_TW"] = str(tw)
            os.environ["B200IMG_RI_BAND"] = str(b)
            row.append(f"{t(call):.4f}")
        print(f"{name} TW={tw:3d} bands {bands}: " + " ".join(row), flush=True)
    del src, dst
