"""Does the library's kernel choice ever lose to the table-driven kernels?  Times default against RESIZE_GENERAL over a spread of
integer-ratio shapes and prints the ratio (> 1.05 = the fixed-window kernel should have declined).   python tools/check_choice.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from image_b200 import FilterType  # noqa: E402
from image_b200 import _lib as L  # noqa: E402

lib = L.lib()
st = torch.cuda.current_stream().cuda_stream


def t(fn, n=30):
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


DT = {"u8": (torch.uint8, L.U8), "u16": (torch.uint16, L.U16), "f32": (torch.float32, L.F32)}
cases = []
for f in (FilterType.Triangle, FilterType.CatmullRom, FilterType.Lanczos3):
    for (n, w, h) in ((1, 640, 480), (1, 1920, 1080), (1, 3840, 2160), (1, 256, 4096), (1, 8192, 128), (8, 512, 512), (64, 256, 256), (2, 4096, 4096)):
        for r in (2, 3, 4):
            for c, d in ((4, "u8"), (3, "u8"), (1, "u8"), (4, "u16"), (4, "f32")):
                if r == 3 and c != 4:
                    continue
                if w % r or h % r or (w * c) % 4:
                    continue
                cases.append((f, n, w, h, r, c, d, False))
    for (n, w, h) in ((1, 640, 480), (1, 1920, 1080), (4, 1024, 1024), (1, 4096, 4096), (16, 512, 512)):
        for d in ("u8", "u16", "f32"):
            cases.append((f, n, w, h, 2, 4, d, True))
worst = []
for f, n, w, h, r, c, d, up in cases:
    nw, nh = (w * 2, h * 2) if up else (w // r, h // r)
    src = (torch.rand((n, h, w, c), device="cuda") * (1 if d == "f32" else 255 if d == "u8" else 65535)).to(DT[d][0])
    dst = torch.empty((n, nh, nw, c), dtype=DT[d][0], device="cuda")
    call = lambda fl: L.check(lib.b200img_resize_batch_dev(src.data_ptr(), 0, w, h, c, DT[d][1], dst.data_ptr(), 0, nw, nh, n, int(f), 1.0, fl, st))
    a = t(lambda: call(L.FP_FAST))
    b = t(lambda: call(L.FP_FAST | L.RESIZE_GENERAL))
    tag = f"{f.name:10s} {n}x{w}x{h}x{c} {d} {'up2' if up else '/' + str(r)}"
    worst.append((a / b, tag, a, b))
worst.sort(reverse=True)
for ratio, tag, a, b in worst[:25]:
    print(f"{ratio:5.2f}  {tag:40s} default {a:8.1f} us  general {b:8.1f} us")
print("...")
for ratio, tag, a, b in worst[-8:]:
    print(f"{ratio:5.2f}  {tag:40s} default {a:8.1f} us  general {b:8.1f} us")
