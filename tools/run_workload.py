"""Runs one hot-path workload a few times (for ncu captures).  usage: run_workload.py NAME [reps]"""
import ctypes as C
import sys, os
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from image_b200 import FilterType, GaussianBlurParameters
from image_b200 import _lib as L

name = sys.argv[1]
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
dev = torch.device("cuda:0")
lib = L.lib()
g = torch.Generator(device=dev).manual_seed(1)
st = torch.cuda.current_stream().cuda_stream
if name in ("k4", "k1", "k5"):
    if name == "k4":
        n, w, h, nw, nh, f, dt, tdt = 64, 1024, 1024, 256, 256, FilterType.CatmullRom, L.U8, torch.uint8
    elif name == "k1":
        n, w, h, nw, nh, f, dt, tdt = 64, 1920, 1080, 960, 540, FilterType.Lanczos3, L.U8, torch.uint8
    else:
        n, w, h, nw, nh, f, dt, tdt = 1, 4096, 2048, 8192, 4096, FilterType.Lanczos3, L.F32, torch.float32
    src = (torch.rand((n, h, w, 4), device=dev, generator=g) * (255 if tdt == torch.uint8 else 1)).to(tdt)
    dst = torch.empty((n, nh, nw, 4), dtype=tdt, device=dev)
    fn = lambda: L.check(lib.b200img_resize_batch_dev(src.data_ptr(), 0, w, h, 4, dt, dst.data_ptr(), 0, nw, nh, n, int(f), 1.0, 0, st))
elif name in ("gauss65", "gauss21", "gauss65q", "fast", "gauss65x8", "fastx8"):
    n = 8 if name.endswith("x8") else 2
    name = name[:-2] if name.endswith("x8") else name
    src = torch.randint(0, 256, (n, 2160, 3840, 3), dtype=torch.uint8, device=dev, generator=g)
    dst = torch.empty_like(src)
    if name == "fast":
        fn = lambda: L.check(lib.b200img_fast_blur_batch_dev(src.data_ptr(), 3840, 2160, 3, L.U8, dst.data_ptr(), n, 10.0, st))
    else:
        p = GaussianBlurParameters.new_from_radius(10.0) if name == "gauss21" else GaussianBlurParameters.new_from_sigma(10.0)
        path = L.BLUR_DYN if name.endswith("q") else L.BLUR_GENERIC
        fn = lambda: L.check(lib.b200img_gaussian_blur_batch_dev(src.data_ptr(), 3840, 2160, 3, L.U8, dst.data_ptr(), n, C.byref(p.raw), path, 0, st))
elif name == "k3":
    npx = 8192 * 8192
    a16 = torch.randint(0, 65536, (npx, 4), dtype=torch.int32, device=dev, generator=g).to(torch.uint16)
    o8 = torch.empty((npx, 3), dtype=torch.uint8, device=dev)
    a8 = torch.randint(0, 256, (npx, 3), dtype=torch.uint8, device=dev, generator=g)
    of = torch.empty((npx, 4), dtype=torch.float32, device=dev)
    bg = torch.randint(0, 256, (npx, 4), dtype=torch.uint8, device=dev, generator=g)

    def fn():
        L.check(lib.b200img_convert_dev(a16.data_ptr(), L.U16, 4, o8.data_ptr(), L.U8, 3, npx, L.CAST, None, st))
        L.check(lib.b200img_convert_dev(a8.data_ptr(), L.U8, 3, of.data_ptr(), L.F32, 4, npx, L.CAST, None, st))
        L.check(lib.b200img_swizzle_02_dev(bg.data_ptr(), L.U8, 4, npx * 4, st))
else:
    raise SystemExit(f"unknown workload {name}")
for _ in range(reps):
    fn()
torch.cuda.synchronize()
print("ok", name)
