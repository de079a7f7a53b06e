"""Host-side mirror of the reference's `ImageBuffer` / `DynamicImage` methods on the hot path.

    ImageBuffer::convert                 src/images/buffer.rs:1643-1654
    ImageBuffer::{from,into}_raw_bgr(a)  src/images/buffer.rs:1544-1582
    DynamicImage::to_{luma,luma_alpha,rgb,rgba}{8,16,32f}   src/images/dynimage.rs:320-424
    DynamicImage::{resize,resize_exact,resize_to_fill}      src/images/dynimage.rs:1127-1206
    DynamicImage::{blur,blur_advanced,fast_blur}            src/images/dynimage.rs:1226-1261

Pixel types are spelled as in the reference's ColorType: "L8", "La8", "Rgb8", "Rgba8",
"L16", ..., "Rgb32F", "Rgba32F" (Luma/LumaA 32F do not exist as DynamicImage variants but
are accepted by ImageBuffer).
"""
from __future__ import annotations

import numpy as np

from . import _lib as L
from . import imageops
from .imageops import FilterType, GaussianBlurParameters, PanicError

_LAYOUT = {"L": 1, "La": 2, "Rgb": 3, "Rgba": 4}
_DEPTH = {"8": np.uint8, "16": np.uint16, "32F": np.float32}


def parse_pixel(name: str):
    for lay in ("Rgba", "Rgb", "La", "L"):
        if name.startswith(lay) and name[len(lay):] in _DEPTH:
            return _LAYOUT[lay], np.dtype(_DEPTH[name[len(lay):]])
    raise ValueError(f"unknown pixel type {name!r}")


def pixel_name(channels: int, dtype) -> str:
    lay = {v: k for k, v in _LAYOUT.items()}[channels]
    dep = {np.dtype(v): k for k, v in _DEPTH.items()}[np.dtype(dtype)]
    return lay + dep


class ImageBuffer:
    """`ImageBuffer<P, Vec<S>>`: `data` is a (h, w, c) numpy array (host) or CUDA tensor (device-resident)."""

    def __init__(self, data):
        if data.ndim == 2:
            data = data[:, :, None]
        if data.ndim != 3 or not 1 <= data.shape[2] <= 4:
            raise ValueError(f"expected (h, w, c<=4), got {tuple(data.shape)}")
        self.data = data

    @property
    def width(self): return int(self.data.shape[1])
    @property
    def height(self): return int(self.data.shape[0])
    @property
    def channels(self): return int(self.data.shape[2])
    def dimensions(self): return self.width, self.height

    def _np_dtype(self):
        if imageops._is_torch(self.data):
            import torch
            return np.dtype({torch.uint8: np.uint8, torch.uint16: np.uint16, torch.float32: np.float32}[self.data.dtype])
        return self.data.dtype

    @property
    def pixel(self) -> str:
        return pixel_name(self.channels, self._np_dtype())

    def convert(self, to_pixel: str) -> "ImageBuffer":
        """ImageBuffer::convert::<To>() -- FromColor arithmetic (integer sRGB luma)."""
        c, dt = parse_pixel(to_pixel)
        return ImageBuffer(imageops.convert_pixels(self.data, c, dt, mode=L.CONVERT))

    @staticmethod
    def from_raw_bgr(width, height, buf): return ImageBuffer._from_raw_swz(width, height, buf, 3)
    @staticmethod
    def from_raw_bgra(width, height, buf): return ImageBuffer._from_raw_swz(width, height, buf, 4)

    @staticmethod
    def _from_raw_swz(width, height, buf, c):
        """from_raw_bgr(a): takes ownership of `buf` and swizzles it in place; None if too small."""
        n = width * height * c
        flat = buf.reshape(-1)
        if flat.shape[0] < n:
            return None
        flat = flat[:n]
        imageops.swizzle_02(flat, c)
        return ImageBuffer(flat.reshape(height, width, c))

    def into_raw_bgr(self): return self._into_raw_swz(3)
    def into_raw_bgra(self): return self._into_raw_swz(4)

    def _into_raw_swz(self, c):
        if self.channels != c:
            raise TypeError(f"into_raw_bg{'ra' if c == 4 else 'r'} exists only for {c}-channel buffers")
        flat = self.data.reshape(-1)
        imageops.swizzle_02(flat, c)
        return flat


class DynamicImage:
    """The DynamicImage methods on the hot path, over one ImageBuffer variant."""

    VARIANTS = ("L8", "La8", "Rgb8", "Rgba8", "L16", "La16", "Rgb16", "Rgba16", "Rgb32F", "Rgba32F")

    def __init__(self, buf: ImageBuffer, luma_coefs=None):
        if buf.pixel not in self.VARIANTS:
            raise ValueError(f"{buf.pixel} is not a DynamicImage variant")
        self.buf = buf
        # CicpRgb::derived_luminance of the image's primaries (src/metadata/cicp.rs:1240-1277);
        # None = sRGB fallback [0.2126, 0.7152, 0.0722] (src/images/buffer.rs:1743)
        self.luma_coefs = luma_coefs

    @property
    def width(self): return self.buf.width
    @property
    def height(self): return self.buf.height
    def color(self) -> str: return self.buf.pixel

    def _to(self, name: str) -> ImageBuffer:
        if self.buf.pixel == name:  # same variant: clone
            d = self.buf.data
            return ImageBuffer(d.clone() if imageops._is_torch(d) else d.copy())
        c, dt = parse_pixel(name)
        return ImageBuffer(imageops.convert_pixels(self.buf.data, c, dt, mode=L.CAST, coefs=self.luma_coefs))

    def to_luma8(self): return self._to("L8")
    def to_luma_alpha8(self): return self._to("La8")
    def to_rgb8(self): return self._to("Rgb8")
    def to_rgba8(self): return self._to("Rgba8")
    def to_luma16(self): return self._to("L16")
    def to_luma_alpha16(self): return self._to("La16")
    def to_rgb16(self): return self._to("Rgb16")
    def to_rgba16(self): return self._to("Rgba16")
    def to_rgb32f(self): return self._to("Rgb32F")
    def to_rgba32f(self): return self._to("Rgba32F")

    # The reference routes DynamicImage::resize* to the third-party pic-scale-safe crate
    # (src/imageops/resize.rs); per the project's north star the drop-in runs the in-tree
    # sample.rs arithmetic instead ("resize boundary").
    # The control flow around it is the reference's (resize.rs:13-93): same size -> unchanged; any
    # filter but Nearest on an image whose alpha is not constant -> premultiply, resample,
    # unpremultiply (that arithmetic is pic-scale-safe's too: parity unpinned; `premultiply=False`
    # resamples straight alpha).
    def resize_exact(self, nwidth, nheight, filter: FilterType, *, premultiply: bool = True, **kw) -> "DynamicImage":
        if (nwidth, nheight) == (self.width, self.height):
            return self
        d = self.buf.data
        pre = premultiply and filter != FilterType.Nearest and not imageops.has_constant_alpha(d)
        if pre:
            d = d.clone() if imageops._is_torch(d) else d.copy()
            imageops.premultiply_alpha(d)
        out = imageops.resize(d, nwidth, nheight, filter, **kw)
        if pre:
            imageops.unpremultiply_alpha(out)
        return DynamicImage(ImageBuffer(out), self.luma_coefs)

    def resize(self, nwidth, nheight, filter: FilterType, **kw) -> "DynamicImage":
        if (nwidth, nheight) == (self.width, self.height):
            return self
        w, h = imageops.resize_dimensions(self.width, self.height, nwidth, nheight, False)
        return self.resize_exact(w, h, filter, **kw)

    def resize_to_fill(self, nwidth, nheight, filter: FilterType, **kw) -> "DynamicImage":
        w2, h2 = imageops.resize_dimensions(self.width, self.height, nwidth, nheight, True)
        inter = self.resize_exact(w2, h2, filter, **kw)
        iw, ih = inter.width, inter.height
        d = inter.buf.data
        # crop selection, src/images/dynimage.rs:1193-1203
        if nwidth * ih > iw * nheight:
            y0 = (ih - nheight) // 2
            d = d[y0:y0 + nheight, 0:nwidth]
        else:
            x0 = (iw - nwidth) // 2
            d = d[0:nheight, x0:x0 + nwidth]
        d = d.contiguous() if imageops._is_torch(d) else np.ascontiguousarray(d)
        return DynamicImage(ImageBuffer(d), self.luma_coefs)

    def blur(self, radius: float) -> "DynamicImage":
        return self.blur_advanced(GaussianBlurParameters.new_from_radius(radius))

    def blur_advanced(self, parameters: GaussianBlurParameters) -> "DynamicImage":
        return DynamicImage(ImageBuffer(imageops.blur_dyn(self.buf.data, parameters)), self.luma_coefs)

    def fast_blur(self, sigma: float) -> "DynamicImage":
        return DynamicImage(ImageBuffer(imageops.fast_blur(self.buf.data, sigma)), self.luma_coefs)

    def unsharpen(self, sigma: float, threshold: int) -> "DynamicImage":
        """DynamicImage::unsharpen (src/images/dynimage.rs:1277): imageops::unsharpen per variant."""
        return DynamicImage(ImageBuffer(imageops.unsharpen(self.buf.data, sigma, threshold)), self.luma_coefs)

    def filter3x3(self, kernel) -> "DynamicImage":
        """DynamicImage::filter3x3 (src/images/dynimage.rs:1291)."""
        return DynamicImage(ImageBuffer(imageops.filter3x3(self.buf.data, kernel)), self.luma_coefs)

    def thumbnail_exact(self, nwidth: int, nheight: int) -> "DynamicImage":
        """DynamicImage::thumbnail_exact (src/images/dynimage.rs:1175)."""
        return DynamicImage(ImageBuffer(imageops.thumbnail(self.buf.data, nwidth, nheight)), self.luma_coefs)

    def thumbnail(self, nwidth: int, nheight: int) -> "DynamicImage":
        """DynamicImage::thumbnail (src/images/dynimage.rs:1159): aspect ratio kept, no fill."""
        w2, h2 = imageops.resize_dimensions(self.width, self.height, nwidth, nheight, False)
        return self.thumbnail_exact(w2, h2)

    # ---- colorops (src/images/dynimage.rs:1082-1118, 1302-1339) ----
    def grayscale(self) -> "DynamicImage":
        """Luma stays Luma (clone); LumaA / Rgba keep their alpha (grayscale_alpha); Rgb -> Luma."""
        c = self.buf.channels
        if c == 1:
            d = self.buf.data
            return DynamicImage(ImageBuffer(d.clone() if imageops._is_torch(d) else d.copy()), self.luma_coefs)
        return DynamicImage(ImageBuffer(imageops.grayscale(self.buf.data, alpha=c in (2, 4))), self.luma_coefs)

    def invert(self) -> None:
        """In place, like DynamicImage::invert(&mut self)."""
        imageops.invert(self.buf.data)

    def adjust_contrast(self, c: float) -> "DynamicImage":
        return DynamicImage(ImageBuffer(imageops.contrast(self.buf.data, c)), self.luma_coefs)

    def brighten(self, value: int) -> "DynamicImage":
        return DynamicImage(ImageBuffer(imageops.brighten(self.buf.data, value)), self.luma_coefs)

    def huerotate(self, value: int) -> "DynamicImage":
        return DynamicImage(ImageBuffer(imageops.huerotate(self.buf.data, value)), self.luma_coefs)
