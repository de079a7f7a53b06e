"""Times the Gaussian blur kernels on the bench's main shape (8 x 3840x2160 Rgb8, sigma 10) and a few others:
fused kernel (with forced CTA counts) against the row / column kernel pair.   python tools/time_gauss.py"""
import ctypes as C
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from image_b200 import GaussianBlurParameters  # noqa: E402
from image_b200 import _lib as L  # noqa: E402

lib = L.lib()
dev = torch.device("cuda:0")
stream = torch.cuda.current_stream().cuda_stream


def t(fn, n=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def run(name, shape, dtype, sigma, path, bands_list):
    n, h, w, c = shape
    if dtype == torch.float32:
        src = torch.rand(shape, device=dev)
    else:
        src = torch.randint(0, 256 if dtype == torch.uint8 else 65536, shape, device=dev, dtype=torch.int32).to(dtype)
    dst = torch.empty_like(src)
    dt = {torch.uint8: L.U8, torch.uint16: L.U16, torch.float32: L.F32}[dtype]
    p = GaussianBlurParameters.new_from_sigma(sigma)
    res = {}
    for flags, label in ((L.FP_FAST | L.GAUSS_TWO_KERNELS, "two-kernel"), (L.FP_FAST | L.GAUSS_FUSED, "fused"), (L.FP_FAST, "auto")):
        for b in (bands_list if label == "fused" else [0]):
            if b:
                os.environ["B200IMG_GF_CTAS"] = str(b)
            else:
                os.environ.pop("B200IMG_GF_CTAS", None)
            ms = t(lambda: L.check(lib.b200img_gaussian_blur_batch_dev(src.data_ptr(), w, h, c, dt, dst.data_ptr(), n, C.byref(p.raw), path, flags, stream)))
            res[f"{label}{'/ctas' + str(b) if b else ''}"] = round(ms, 4)
    os.environ.pop("B200IMG_GF_CTAS", None)
    print(name, res, flush=True)


if __name__ == "__main__":
    run("8x4K rgb8 sigma10 generic", (8, 2160, 3840, 3), torch.uint8, 10.0, L.BLUR_GENERIC, [0, 74, 148, 296])
    run("8x4K rgb8 sigma10 q15", (8, 2160, 3840, 3), torch.uint8, 10.0, L.BLUR_DYN, [0])
    run("8x4K rgb8 radius10(sigma3.5) generic", (8, 2160, 3840, 3), torch.uint8, 3.5, L.BLUR_GENERIC, [0])
    run("8x4K rgba8 sigma10", (8, 2160, 3840, 4), torch.uint8, 10.0, L.BLUR_GENERIC, [0])
    run("8x4K luma8 sigma10", (8, 2160, 3840, 1), torch.uint8, 10.0, L.BLUR_GENERIC, [0])
    run("4x4K rgb16 sigma10", (4, 2160, 3840, 3), torch.uint16, 10.0, L.BLUR_GENERIC, [0])
    run("2x4K rgb32f sigma10", (2, 2160, 3840, 3), torch.float32, 10.0, L.BLUR_GENERIC, [0])
    run("1x4K rgb8 sigma10", (1, 2160, 3840, 3), torch.uint8, 10.0, L.BLUR_GENERIC, [0, 37, 74])
    run("1x1024x768 rgb8 sigma3", (1, 768, 1024, 3), torch.uint8, 3.0, L.BLUR_GENERIC, [0, 24, 48, 96])
    run("1x1024x768 rgb8 sigma7", (1, 768, 1024, 3), torch.uint8, 7.0, L.BLUR_GENERIC, [0, 24, 48, 96])
    run("16x512x512 rgba8 sigma2", (16, 512, 512, 4), torch.uint8, 2.0, L.BLUR_GENERIC, [0, 64])
