"""End-to-end (host-pointer) Gaussian blur, sigma 10, on n 4K Rgb8 frames in pinned memory, for several n:
per-frame time against batch size shows how much of the call is pipeline fill and drain."""
import ctypes as C, os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from image_b200 import GaussianBlurParameters
from image_b200 import _lib as L
lib = L.lib()
torch.cuda.set_device(0)
p = GaussianBlurParameters.new_from_sigma(10.0)
w, h = 3840, 2160
for n in (1, 2, 4, 8, 16, 32):
    src = torch.randint(0, 256, (n, h, w, 3), dtype=torch.uint8).pin_memory()
    dst = torch.empty_like(src).pin_memory()
    call = lambda: L.check(lib.b200img_gaussian_blur_batch(src.data_ptr(), w, h, 3, L.U8, dst.data_ptr(), n, C.byref(p.raw), L.BLUR_GENERIC, 0))
    for _ in range(3):
        call()
    t0 = time.perf_counter()
    reps = 10
    for _ in range(reps):
        call()
    ms = (time.perf_counter() - t0) / reps * 1e3
    print(f"n={n:3d}  {ms:8.3f} ms per call  {ms / n:7.3f} ms per frame  {n * w * h / 1e6 / (ms / 1e3) / 1e3:7.2f} GP/s")
