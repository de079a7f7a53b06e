"""Times the resampler at the BASELINE shapes K1 / K4 and a few neighbours.   python tools/time_resize.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from image_b200 import FilterType  # noqa: E402
from image_b200 import _lib as L  # noqa: E402

lib = L.lib()
dev = torch.device("cuda:0")
st = torch.cuda.current_stream().cuda_stream
hbm = 6551.0


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


CASES = [("K1 lanczos3 64x1080p rgba8 -> 960x540", 64, 1920, 1080, 4, torch.uint8, 960, 540, FilterType.Lanczos3),
         ("K4 catmullrom 256x1024^2 rgba8 -> 256^2", 256, 1024, 1024, 4, torch.uint8, 256, 256, FilterType.CatmullRom),
         ("lanczos3 64x1080p rgb8 -> 960x540", 64, 1920, 1080, 3, torch.uint8, 960, 540, FilterType.Lanczos3),
         ("triangle 64x1080p rgba8 -> 640x360", 64, 1920, 1080, 4, torch.uint8, 640, 360, FilterType.Triangle),
         ("lanczos3 16x4K rgba8 -> 1280x720", 16, 3840, 2160, 4, torch.uint8, 1280, 720, FilterType.Lanczos3),
         ("lanczos3 32x1080p rgba16 -> 960x540", 32, 1920, 1080, 4, torch.uint16, 960, 540, FilterType.Lanczos3),
         ("lanczos3 16x1080p rgba32f -> 960x540", 16, 1920, 1080, 4, torch.float32, 960, 540, FilterType.Lanczos3),
         ("gaussian 64x1080p luma8 -> 960x540", 64, 1920, 1080, 1, torch.uint8, 960, 540, FilterType.Gaussian),
         ("catmullrom 16x4K rgb8 -> 1280x720", 16, 3840, 2160, 3, torch.uint8, 1280, 720, FilterType.CatmullRom)]
for name, n, w, h, c, dt, nw, nh, f in CASES:
    src = torch.rand((n, h, w, c), device=dev) if dt == torch.float32 else torch.randint(0, 256 if dt == torch.uint8 else 65536, (n, h, w, c), dtype=torch.int32, device=dev).to(dt)
    dst = torch.empty((n, nh, nw, c), dtype=dt, device=dev)
    ldt = L.U8 if dt == torch.uint8 else L.U16 if dt == torch.uint16 else L.F32
    ms = t(lambda: L.check(lib.b200img_resize_batch_dev(src.data_ptr(), 0, w, h, c, ldt, dst.data_ptr(), 0, nw, nh, n, int(f), 1.0, L.FP_FAST, st)))
    alg = n * (w * h + nw * nh) * c * src.element_size()
    print(f"{name:45s} {ms:.4f} ms  {alg / ms / 1e6:8.1f} GB/s  {alg / ms / 1e6 / hbm:.3f} of HBM", flush=True)
    del src, dst
