"""Timing experiments on the fused Gaussian kernel (results are WRONG with any dbg bit set): which part of the
kernel the time goes to.  python tools/time_gauss_dbg.py"""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from image_b200 import GaussianBlurParameters  # noqa: E402
from image_b200 import _lib as L  # noqa: E402

lib = L.lib()
dev = torch.device("cuda:0")
st = torch.cuda.current_stream().cuda_stream
src = torch.randint(0, 256, (8, 2160, 3840, 3), dtype=torch.uint8, device=dev)
dst = torch.empty_like(src)
p = GaussianBlurParameters.new_from_sigma(10.0)


def t(n=20):
    f = lambda: L.check(lib.b200img_gaussian_blur_batch_dev(src.data_ptr(), 3840, 2160, 3, L.U8, dst.data_ptr(), 8, C.byref(p.raw), L.BLUR_GENERIC, 0, st))
    for _ in range(3):
        f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        f()
    e1.record()
    torch.cuda.synchronize()
    return round(e0.elapsed_time(e1) / n, 4)


for dbg, what in ((0, "everything"), (1, "no staging work"), (2, "no output stores"), (4, "no ring stores"), (3, "no staging, no output stores"),
                  (7, "only barriers + multiply-adds"), (8, "no multiply-adds"), (15, "barriers only")):
    os.environ["B200IMG_GF_DBG"] = str(dbg)
    print(f"dbg {dbg:2d} {what:32s} {t()} ms", flush=True)
