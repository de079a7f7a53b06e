"""Times fast_blur (sigma 10, u8) fused against the horizontal + vertical kernel pair, with forced band counts.
python tools/time_fast_blur.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from image_b200 import _lib as L  # noqa: E402

lib = L.lib()
dev = torch.device("cuda:0")
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
    return round(e0.elapsed_time(e1) / n, 4)


for name, shape, sigma in (("8x4K rgb8 sigma10", (8, 2160, 3840, 3), 10.0), ("8x4K rgba8 sigma10", (8, 2160, 3840, 4), 10.0),
                           ("8x4K luma8 sigma10", (8, 2160, 3840, 1), 10.0), ("8x4K rgb8 sigma3", (8, 2160, 3840, 3), 3.0),
                           ("1x4K rgb8 sigma10", (1, 2160, 3840, 3), 10.0), ("1x1024x768 rgb8 sigma7", (1, 768, 1024, 3), 7.0),
                           ("64x512x512 rgba8 sigma4", (64, 512, 512, 4), 4.0)):
    n, h, w, c = shape
    src = torch.randint(0, 256, shape, dtype=torch.uint8, device=dev)
    dst = torch.empty_like(src)
    fn = lambda: L.check(lib.b200img_fast_blur_batch_dev(src.data_ptr(), w, h, c, L.U8, dst.data_ptr(), n, sigma, st))
    res = {}
    os.environ["B200IMG_BOX_UNFUSED"] = "1"
    res["kernel pair"] = t(fn)
    del os.environ["B200IMG_BOX_UNFUSED"]
    for b in (0, 1, 2, 3, 4, 6, 8, 16):
        if b:
            os.environ["B200IMG_BOX_BANDS"] = str(b)
        else:
            os.environ.pop("B200IMG_BOX_BANDS", None)
        res[f"fused{'/b' + str(b) if b else ''}"] = t(fn)
    os.environ.pop("B200IMG_BOX_BANDS", None)
    print(name, res, flush=True)
