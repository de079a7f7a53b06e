"""One-off stress of the conversion / swizzle kernels: random layouts, sizes (tails) and buffer offsets (alignment fallbacks)
against the oracle.   python tools/stress_convert.py [cases] [seed]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import oracle as o  # noqa: E402
from image_b200 import _lib as L  # noqa: E402

lib = L.lib()
st = torch.cuda.current_stream().cuda_stream
cases = int(sys.argv[1]) if len(sys.argv) > 1 else 300
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 3)
DT = [(np.uint8, L.U8, torch.uint8), (np.uint16, L.U16, torch.uint16), (np.float32, L.F32, torch.float32)]
bad = 0
for it in range(cases):
    sd, dd = DT[int(rng.integers(0, 3))], DT[int(rng.integers(0, 3))]
    sc, dc = int(rng.integers(1, 5)), int(rng.integers(1, 5))
    mode = [L.CAST, L.CONVERT][int(rng.integers(0, 2))]
    n = int(rng.choice([1, 37, 4095, 4096, 4097, 5000 + int(rng.integers(0, 300)), 70001, 262144 + int(rng.integers(0, 200))]))
    so, do = int(rng.integers(0, 5)), int(rng.integers(0, 5))   # offsets in elements: 0 keeps 16-byte alignment
    if sd[0] == np.float32:
        a = rng.random((n, sc), dtype=np.float32) * 1.2 - 0.1
    else:
        a = rng.integers(0, np.iinfo(sd[0]).max + 1, (n, sc), dtype=sd[0])
    sbuf = torch.zeros(n * sc + 8, dtype=sd[2], device="cuda")
    sbuf[so:so + n * sc] = torch.from_numpy(a.reshape(-1)).cuda()
    dbuf = torch.zeros(n * dc + 8, dtype=dd[2], device="cuda")
    L.check(lib.b200img_convert_dev(sbuf.data_ptr() + so * sbuf.element_size(), sd[1], sc, dbuf.data_ptr() + do * dbuf.element_size(), dd[1], dc, n,
                                    mode, None, st))
    got = dbuf[do:do + n * dc].cpu().numpy().reshape(n, dc)
    want = o.convert(a, sc, dd[0], dc, mode)
    if got.tobytes() != np.ascontiguousarray(want).tobytes() and not (dd[0] == np.float32 and np.array_equal(got, want)):
        bad += 1
        print("MISMATCH convert", sd[0].__name__, sc, "->", dd[0].__name__, dc, "mode", mode, "n", n, "offsets", so, do, flush=True)
    if bool((dbuf[:do].cpu().numpy() != 0).any()) or bool((dbuf[do + n * dc:].cpu().numpy() != 0).any()):
        bad += 1
        print("WROTE OUTSIDE convert", sd[0].__name__, sc, "->", dd[0].__name__, dc, n, so, do, flush=True)
    # swizzle in place
    ch = int(rng.choice([3, 4]))
    if sd[0] == np.float32:
        b = rng.random(n * ch + int(rng.integers(0, ch)), dtype=np.float32)
    else:
        b = rng.integers(0, np.iinfo(sd[0]).max + 1, n * ch + int(rng.integers(0, ch)), dtype=sd[0])
    wbuf = torch.zeros(b.size + 8, dtype=sd[2], device="cuda")
    wbuf[so:so + b.size] = torch.from_numpy(b).cuda()
    L.check(lib.b200img_swizzle_02_dev(wbuf.data_ptr() + so * wbuf.element_size(), sd[1], ch, b.size, st))
    if wbuf[so:so + b.size].cpu().numpy().tobytes() != o.swizzle_02(b, ch).tobytes() or bool((wbuf[:so].cpu().numpy() != 0).any()) or bool((wbuf[so + b.size:].cpu().numpy() != 0).any()):
        bad += 1
        print("MISMATCH swizzle", sd[0].__name__, ch, b.size, so, flush=True)
print(f"{cases} cases, {bad} problems")
