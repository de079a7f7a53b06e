"""The K3 conversions (and a few neighbours) through the tile kernel and the direct kernel.   python tools/time_convert.py"""
import os
import subprocess
import sys

if len(sys.argv) == 1:
    for d in ("0", "1"):
        subprocess.run([sys.executable, __file__, d], env=dict(os.environ, B200IMG_CONVERT_DIRECT=d))
    sys.exit(0)

import torch  # noqa: E402

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
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


npx = 8192 * 8192
DT = {"u8": (torch.uint8, L.U8, 1), "u16": (torch.uint16, L.U16, 2), "f32": (torch.float32, L.F32, 4)}
print("direct" if sys.argv[1] == "1" else "tile kernel")
for sd, ci, dd, co, mode in (("u16", 4, "u8", 3, L.CAST), ("u8", 3, "f32", 4, L.CAST), ("u8", 4, "u8", 3, L.CAST), ("u8", 3, "u8", 4, L.CAST),
                             ("u8", 4, "f32", 4, L.CAST), ("f32", 4, "u8", 4, L.CAST), ("u8", 3, "u8", 1, L.CONVERT), ("u16", 3, "u16", 4, L.CAST),
                             ("u8", 1, "u8", 4, L.CAST), ("f32", 3, "u16", 3, L.CAST)):
    a = torch.zeros((npx, ci), dtype=DT[sd][0], device="cuda")
    b = torch.empty((npx, co), dtype=DT[dd][0], device="cuda")
    ms = t(lambda: L.check(lib.b200img_convert_dev(a.data_ptr(), DT[sd][1], ci, b.data_ptr(), DT[dd][1], co, npx, mode, None, st)))
    by = npx * (ci * DT[sd][2] + co * DT[dd][2])
    print(f"  {sd}x{ci} -> {dd}x{co}: {ms:.4f} ms {by / ms / 1e6:.0f} GB/s {by / ms / 1e6 / 6551:.3f}", flush=True)
    del a, b
