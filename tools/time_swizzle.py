import sys, torch
sys.path.insert(0, '/root/repo')
from image_b200 import _lib as L
lib = L.lib(); st = torch.cuda.current_stream().cuda_stream
def t(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
npx = 8192 * 8192
for dt, ldt, es in ((torch.uint8, L.U8, 1), (torch.uint16, L.U16, 2), (torch.float32, L.F32, 4)):
    for c in (4, 3):
        buf = torch.zeros((npx, c), dtype=dt, device="cuda")
        ms = t(lambda: L.check(lib.b200img_swizzle_02_dev(buf.data_ptr(), ldt, c, npx * c, st)))
        print(dt, c, f"{ms:.4f} ms {2 * npx * c * es / ms / 1e6:.0f} GB/s {2 * npx * c * es / ms / 1e6 / 6551:.3f}")
        del buf
