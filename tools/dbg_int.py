"""Debug aid: where does the integer-ratio resampler differ from the oracle?   python tools/dbg_int.py"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import oracle as o  # noqa: E402
from image_b200 import FilterType  # noqa: E402
from image_b200 import _lib as L  # noqa: E402

lib = L.lib()
st = torch.cuda.current_stream().cuda_stream
rng = np.random.default_rng(1)
for (n, w, h, c, r, f) in [(24, 512, 384, 4, 4, FilterType.CatmullRom), (24, 512, 384, 4, 2, FilterType.CatmullRom), (24, 512, 384, 3, 4, FilterType.CatmullRom),
                           (24, 768, 384, 4, 4, FilterType.CatmullRom), (24, 512, 384, 4, 4, FilterType.Triangle)]:
    imgs = rng.integers(0, 256, (n, h, w, c), dtype=np.uint8)
    src = torch.from_numpy(imgs).cuda()
    nw, nh = w // r, h // r
    dst = torch.zeros((n, nh, nw, c), dtype=torch.uint8, device="cuda")
    L.check(lib.b200img_resize_batch_dev(src.data_ptr(), 0, w, h, c, L.U8, dst.data_ptr(), 0, nw, nh, n, int(f), 1.0, L.FP_FAST, st))
    got = dst.cpu().numpy().astype(np.int64)
    for i in (0, n - 1):
        want = o.resize(imgs[i], nw, nh, int(f)).astype(np.int64)
        d = np.abs(got[i] - want)
        bad = d > 1
        print((n, w, h, c, r, f.name), "img", i, "max", d.max(), "bad frac", bad.mean(),
              "bad rows", np.unique(np.argwhere(bad)[:, 0])[:12].tolist(), "bad cols", np.unique(np.argwhere(bad)[:, 1])[:40].tolist())
