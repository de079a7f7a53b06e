"""image_b200 -- B200-native (sm_100a) implementation of image-rs/image's pixel hot path.

Resampler (`imageops.resize`), Gaussian blur (`imageops.blur`, `blur_advanced`),
box-approximated blur (`imageops.fast_blur`), pixel-format conversions
(`DynamicImage.to_*`, `ImageBuffer.convert`) and the BGR(A) swizzle, all running
as hand-written CUDA kernels behind the C ABI in `include/b200img.h`.
The package holds the kernels (`csrc/`), the built library (`lib/`), and this
host-side mirror of the reference's interface.  It has no CPU fallback.
"""
from . import _lib, imageops
from ._lib import B200ImgError, GaussParams
from .imageops import FilterType, GaussianBlurParameters, LimitError, PanicError
from .images import DynamicImage, ImageBuffer

__all__ = ["imageops", "FilterType", "GaussianBlurParameters", "ImageBuffer", "DynamicImage",
           "PanicError", "LimitError", "B200ImgError", "GaussParams"]
