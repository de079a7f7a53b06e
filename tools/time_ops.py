"""CUDA-event timings of the section-8f operators (thumbnail, filter3x3, unsharpen) on 8 x 4K Rgb8.
usage: time_ops.py [reps [label-regex]]   -- prints one line per case: ms per call and algorithmic GB/s (src + dst bytes)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from image_b200 import imageops
from image_b200 import _lib as L

import re
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 20
only = re.compile(sys.argv[2]) if len(sys.argv) > 2 else None
dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(1)
lib = L.lib()
st = torch.cuda.current_stream().cuda_stream
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timed(fn, nbytes, label):
    if only and not only.search(label):
        return
    for _ in range(3):
        fn()
    tot = 0.0
    for _ in range(reps):
        flush.zero_()                      # 256 MiB write: evicts the 126 MB L2
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); b.synchronize()
        tot += a.elapsed_time(b)
    ms = tot / reps
    print(f"{label:44s} {ms:8.3f} ms  {nbytes / ms / 1e6:8.1f} GB/s")


# ---- resampler cases outside the bench line: f32 upscale (config 5 shape), u8 upscale, f32 downscale ----
from image_b200 import FilterType
def _resize_case(label, n_, w_, h_, c_, tdt, dt, nw_, nh_, filt):
    if only and not only.search(label):
        return
    if tdt == torch.float32:
        s_ = torch.rand((n_, h_, w_, c_), device=dev, generator=g)
    else:
        s_ = torch.randint(0, 256, (n_, h_, w_, c_), device=dev, generator=g).to(tdt)
    d_ = torch.empty((n_, nh_, nw_, c_), dtype=tdt, device=dev)
    for flags, tag in ((0, ""), (L.RESIZE_TILED, " [tile kernel]")):
        timed(lambda: L.check(lib.b200img_resize_batch_dev(s_.data_ptr(), 0, w_, h_, c_, dt, d_.data_ptr(), 0, nw_, nh_, n_, int(filt), 1.0, flags, st)),
              s_.numel() * s_.element_size() + d_.numel() * d_.element_size(), label + tag)
_resize_case("resize f32 Rgba 8192x8192 -> 16384x16384 Lanczos3", 1, 8192, 8192, 4, torch.float32, L.F32, 16384, 16384, FilterType.Lanczos3)
_resize_case("resize u8 Rgb 16x1920x1080 -> 2000x2000 Lanczos3", 16, 1920, 1080, 3, torch.uint8, L.U8, 2000, 2000, FilterType.Lanczos3)
_resize_case("resize u8 Rgba 16x1920x1080 -> 3840x2160 CatmullRom", 16, 1920, 1080, 4, torch.uint8, L.U8, 3840, 2160, FilterType.CatmullRom)
_resize_case("resize f32 Rgba 16x1920x1080 -> 960x540 Lanczos3", 16, 1920, 1080, 4, torch.float32, L.F32, 960, 540, FilterType.Lanczos3)
# the reference's own bench shapes (benches/imageops.rs:95-124), one image per call
for _f in (FilterType.Nearest, FilterType.Triangle, FilterType.CatmullRom, FilterType.Gaussian, FilterType.Lanczos3):
    _resize_case(f"resize u8 Rgb 1x1920x1080 -> 400x300 {_f.name}", 1, 1920, 1080, 3, torch.uint8, L.U8, 400, 300, _f)
    _resize_case(f"resize u8 Rgb 1x1920x1080 -> 2000x2000 {_f.name}", 1, 1920, 1080, 3, torch.uint8, L.U8, 2000, 2000, _f)
_resize_case("resize u16 Rgb 16x1920x1080 -> 2560x1440 Triangle", 16, 1920, 1080, 3, torch.uint16, L.U16, 2560, 1440, FilterType.Triangle)

# ---- the reference's blur bench shapes (benches/blur.rs): one 1024x768 Rgb8 image, sigma 3 / 7 / 50 ----
from image_b200 import GaussianBlurParameters
import ctypes as C
_b = torch.randint(0, 256, (768, 1024, 3), device=dev, generator=g).to(torch.uint8)
_bo = torch.empty_like(_b)
for _s in (3.0, 7.0, 50.0):
    _p = GaussianBlurParameters.new_from_sigma(_s)
    timed(lambda: L.check(lib.b200img_fast_blur_dev(_b.data_ptr(), 1024, 768, 3, L.U8, _bo.data_ptr(), _s, st)), 2 * _b.numel(), f"refbench fast_blur 1024x768 Rgb8 sigma {_s:g}")
    timed(lambda: L.check(lib.b200img_gaussian_blur_dev(_b.data_ptr(), 1024, 768, 3, L.U8, _bo.data_ptr(), C.byref(_p.raw), L.BLUR_GENERIC, 0, st)),
          2 * _b.numel(), f"refbench blur_advanced 1024x768 Rgb8 sigma {_s:g} ({_p.raw.x_kernel_size} taps)")
    timed(lambda: L.check(lib.b200img_gaussian_blur_dev(_b.data_ptr(), 1024, 768, 3, L.U8, _bo.data_ptr(), C.byref(_p.raw), L.BLUR_DYN, 0, st)),
          2 * _b.numel(), f"refbench DynamicImage::blur_advanced 1024x768 Rgb8 sigma {_s:g}")

n, w, h = 8, 3840, 2160
# ---- either side of DynamicImage::resize: Rgba8 ----
rgba = torch.randint(0, 256, (n * h, w, 4), device=dev, generator=g).to(torch.uint8)
timed(lambda: imageops.has_constant_alpha(rgba), rgba.numel(), "has_constant_alpha u8 8x4K Rgba (varying; incl. sync + 4-byte D2H)")
rgba[..., 3] = 255
timed(lambda: imageops.has_constant_alpha(rgba), rgba.numel(), "has_constant_alpha u8 8x4K Rgba (constant)")
timed(lambda: imageops.premultiply_alpha(rgba), 2 * rgba.numel(), "premultiply_alpha u8 8x4K Rgba (in place)")
timed(lambda: imageops.unpremultiply_alpha(rgba), 2 * rgba.numel(), "unpremultiply_alpha u8 8x4K Rgba (in place)")
del rgba
for tdt, dt, name in ((torch.uint8, L.U8, "u8"), (torch.uint16, L.U16, "u16"), (torch.float32, L.F32, "f32")):
    if tdt == torch.float32:
        src = torch.rand((n, h, w, 3), device=dev, generator=g)
    else:
        src = torch.randint(0, 256, (n, h, w, 3), device=dev, generator=g).to(tdt)
    es = src.element_size()
    for nw, nh in ((256, 256), (384, 216), (1920, 1080), (3000, 2000)):
        dst = torch.empty((n, nh, nw, 3), dtype=tdt, device=dev)
        for flags, tag in ((0, ""), (L.THUMBNAIL_PER_PIXEL, " per-pixel")):
            fn = lambda: L.check(lib.b200img_thumbnail_batch_dev(src.data_ptr(), w, h, 3, dt, dst.data_ptr(), nw, nh, n, flags, st))
            timed(fn, src.numel() * es + dst.numel() * es, f"thumbnail {name} 8x4K -> {nw}x{nh}{tag}")
    if name != "u16":
        small = src[:, :540, :960].contiguous()
        dst = torch.empty((n, 2160, 3840, 3), dtype=tdt, device=dev)
        fn = lambda: L.check(lib.b200img_thumbnail_batch_dev(small.data_ptr(), 960, 540, 3, dt, dst.data_ptr(), 3840, 2160, n, 0, st))
        timed(fn, small.numel() * es + dst.numel() * es, f"thumbnail {name} 8x(960x540) -> 4K (upscale)")
    tall = src.view(n * h, w, 3)   # the elementwise operators see the batch as one tall image
    nb = 2 * tall.numel() * es
    timed(lambda: imageops.invert(tall), nb, f"invert {name} 8x4K (in place)")
    timed(lambda: imageops.brighten_in_place(tall, 10), nb, f"brighten {name} 8x4K (in place)")
    timed(lambda: imageops.contrast_in_place(tall, 15.0), nb, f"contrast {name} 8x4K (in place)")
    timed(lambda: imageops.huerotate_in_place(tall, 45), nb, f"huerotate {name} 8x4K (in place)")
    gray = torch.empty((n * h, w, 1), dtype=tdt, device=dev)
    timed(lambda: L.check(lib.b200img_grayscale_dev(tall.data_ptr(), w, n * h, 3, dt, gray.data_ptr(), 0, st)),
          tall.numel() * es + gray.numel() * es, f"grayscale {name} 8x4K")
    del gray
    one = src[0]
    timed(lambda: imageops.filter3x3(one, [1, 1, 1, 1, -8, 1, 1, 1, 1]), 2 * one.numel() * es, f"filter3x3 {name} 4K")
    timed(lambda: imageops.unsharpen(one, 2.0, 0), 2 * one.numel() * es, f"unsharpen {name} 4K sigma=2")
    del src
