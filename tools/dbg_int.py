"""Debug aid: where does the 2x upscale kernel differ from the oracle around non-finite samples?"""
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
n, h, w = 4, 120, 200
for f in (FilterType.Lanczos3, FilterType.Gaussian):
    for acc in (1.0, 0.0):
        imgs = rng.random((n, h, w, 4), dtype=np.float32)
        imgs[0, h // 2, w // 2] = np.inf
        imgs[0, 3, 5, 1] = np.nan
        src = torch.from_numpy(imgs).cuda()
        for flags, name in ((L.FP_FAST, "up2"), (L.FP_FAST | L.RESIZE_GENERAL, "general")):
            dst = torch.empty((n, 2 * h, 2 * w, 4), dtype=torch.float32, device="cuda")
            L.check(lib.b200img_resize_batch_dev(src.data_ptr(), 0, w, h, 4, L.F32, dst.data_ptr(), 0, 2 * w, 2 * h, n, int(f), acc, flags, st))
            got = dst.cpu().numpy()[0]
            want = o.resize(imgs[0], 2 * w, 2 * h, int(f), acc_init=acc)
            bad = ~np.isclose(got, want, rtol=1e-5, atol=1e-6)
            idx = np.argwhere(bad)
            print(f.name, acc, name, "bad", bad.sum(), "rows", np.unique(idx[:, 0])[:20].tolist(), "cols", np.unique(idx[:, 1])[:20].tolist(),
                  "ch", np.unique(idx[:, 2]).tolist())
            if bad.sum():
                y, x, c = idx[0]
                print("   first", (y, x, c), "got", got[y, x, c], "want", want[y, x, c])
