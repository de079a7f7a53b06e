import os, sys, ctypes as C, torch
sys.path.insert(0, "/root/repo")
from image_b200 import GaussianBlurParameters
from image_b200 import _lib as L
dev = torch.device("cuda:0"); lib = L.lib(); st = torch.cuda.current_stream().cuda_stream
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
def timed(fn, label, reps=5):
    for _ in range(2): fn()
    tot = 0
    for _ in range(reps):
        flush.zero_(); a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); b.synchronize(); tot += a.elapsed_time(b)
    print(f"{label:60s} {tot/reps:8.3f} ms")
n, w, h = 8, 3840, 2160
p = GaussianBlurParameters.new_from_sigma(10.0)
for tdt, dt, name in ((torch.uint8, L.U8, "u8"), (torch.uint16, L.U16, "u16"), (torch.float32, L.F32, "f32")):
    for c in (3, 4):
        src = (torch.rand((n, h, w, c), device=dev) * (1 if tdt == torch.float32 else 255)).to(tdt)
        dst = torch.empty_like(src)
        timed(lambda: L.check(lib.b200img_fast_blur_batch_dev(src.data_ptr(), w, h, c, dt, dst.data_ptr(), n, 10.0, st)), f"fast_blur sigma10 {name} C{c} 8x4K")
        timed(lambda: L.check(lib.b200img_gaussian_blur_batch_dev(src.data_ptr(), w, h, c, dt, dst.data_ptr(), n, C.byref(p.raw), L.BLUR_GENERIC, 0, st)), f"blur generic sigma10 {name} C{c} 8x4K")
        timed(lambda: L.check(lib.b200img_gaussian_blur_batch_dev(src.data_ptr(), w, h, c, dt, dst.data_ptr(), n, C.byref(p.raw), L.BLUR_DYN, 0, st)), f"blur dyn sigma10 {name} C{c} 8x4K")
        del src, dst
