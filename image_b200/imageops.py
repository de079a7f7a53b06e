"""Host-side mirror of the reference's `image::imageops` functions for the hot path.

Same names, argument meaning and error behaviour as the reference:

    resize(image, nwidth, nheight, filter)      src/imageops/sample.rs:970
    blur(image, radius)                         src/imageops/sample.rs:1034
    blur_advanced(image, parameters)            src/imageops/sample.rs:1050
    fast_blur(image, sigma)                     src/imageops/fast_blur.rs:111

An *image* is a numpy array of shape (h, w, c) (or (h, w)), dtype uint8 / uint16 /
float32 -- the packed interleaved layout of `ImageBuffer<P, Vec<S>>` -- or a CUDA
torch tensor of the same shape (the device-resident variant: no copies, the
call is enqueued on torch's current stream).  Where the reference panics these
raise `PanicError`; where it returns `ImageError::Limits` they raise
`LimitError`.  Every call runs on the GPU through libb200img's C ABI; there is
no CPU path.
"""
from __future__ import annotations

import ctypes as C
import enum

import numpy as np

from . import _lib as L
from ._lib import GaussParams


class PanicError(RuntimeError):
    """The reference would panic (assert!/unwrap/unimplemented!) on this input."""


class LimitError(RuntimeError):
    """ImageError::Limits(LimitErrorKind::DimensionError) in the reference."""


class FilterType(enum.IntEnum):
    """src/imageops/sample.rs:98-115"""
    Nearest = 0
    Triangle = 1
    CatmullRom = 2
    Gaussian = 3
    Lanczos3 = 4


#: the reference's literal accumulator start (sample.rs:547); 0.0 = the intended filter
ACC_INIT_REFERENCE = 1.0

_NP2DT = {np.dtype(np.uint8): L.U8, np.dtype(np.uint16): L.U16, np.dtype(np.float32): L.F32}


def _raise(e: L.B200ImgError):
    if e.status == L.ERR_BAD_ARG:
        raise PanicError(e.message) from None
    if e.status == L.ERR_DIMENSION:
        raise LimitError(e.message) from None
    raise e


def _call(fn, *args):
    rc = fn(*args)
    if rc != 0:
        _raise(L.B200ImgError(rc, L.last_error()))


class GaussianBlurParameters:
    """src/imageops/sample.rs:1081-1295 (constructors validated by the C library)."""

    def __init__(self, raw: GaussParams):
        self.raw = raw

    @staticmethod
    def _make(fn, *a):
        p = GaussParams()
        _call(fn, *a, C.byref(p))
        return GaussianBlurParameters(p)

    @classmethod
    def new_from_radius(cls, radius: float):
        return cls._make(L.lib().b200img_gauss_params_from_radius, float(radius))

    @classmethod
    def new_from_kernel_size(cls, kernel_size: float):
        return cls._make(L.lib().b200img_gauss_params_from_kernel_size, float(kernel_size))

    @classmethod
    def new_anisotropic_kernel_size(cls, x: float, y: float):
        return cls._make(L.lib().b200img_gauss_params_from_anisotropic, float(x), float(y))

    @classmethod
    def new_from_sigma(cls, sigma: float):
        return cls._make(L.lib().b200img_gauss_params_from_sigma, float(sigma))

    def __repr__(self):
        return repr(self.raw)


GaussianBlurParameters.SMOOTHING_3 = GaussianBlurParameters(GaussParams(3, 0.8, 3, 0.8))
GaussianBlurParameters.SMOOTHING_5 = GaussianBlurParameters(GaussParams(5, 1.1, 5, 1.1))
GaussianBlurParameters.SMOOTHING_7 = GaussianBlurParameters(GaussParams(7, 1.4, 7, 1.4))


# ---- image plumbing -------------------------------------------------------------
def _is_torch(x) -> bool:
    return type(x).__module__.startswith("torch")


class _Img:
    """(pointer, w, h, c, dtype) view of a numpy array or CUDA tensor."""

    def __init__(self, a, inplace: bool = False):
        self.torch = _is_torch(a)
        if inplace:
            # an in-place operator on a non-contiguous view would run on a temporary copy and leave
            # the caller's image untouched without a word
            ok = a.is_contiguous() if self.torch else (isinstance(a, np.ndarray) and a.flags.c_contiguous and a.flags.writeable)
            if not ok:
                raise ValueError("in-place operators need a contiguous (and, on the host, writable) image")
        if self.torch:
            import torch
            if not a.is_cuda:
                raise TypeError("torch images must be CUDA tensors (numpy arrays are the host path)")
            a = a.contiguous()
            dt = {torch.uint8: L.U8, torch.uint16: L.U16, torch.float32: L.F32}.get(a.dtype)
            if dt is None:
                raise TypeError(f"unsupported dtype {a.dtype}")
            self.ptr = a.data_ptr()
        else:
            a = np.ascontiguousarray(a)
            dt = _NP2DT.get(a.dtype)
            if dt is None:
                raise TypeError(f"unsupported dtype {a.dtype}")
            self.ptr = a.ctypes.data
        self.a = a
        self.dt = dt
        shape = tuple(a.shape)
        self.squeeze = len(shape) == 2
        if self.squeeze:
            shape = shape + (1,)
        if len(shape) != 3:
            raise ValueError(f"expected (h, w, c), got {tuple(a.shape)}")
        self.h, self.w, self.c = shape
        if self.c > 4:
            raise PanicError("more than 4 channels is unimplemented in the reference")

    def empty_like(self, h=None, w=None, c=None, dtype=None):
        h = self.h if h is None else h
        w = self.w if w is None else w
        c = self.c if c is None else c
        shape = (h, w) if (self.squeeze and c == 1) else (h, w, c)
        if self.torch:
            import torch
            return torch.empty(shape, dtype=dtype or self.a.dtype, device=self.a.device)
        return np.empty(shape, dtype=dtype or self.a.dtype)


def _ptr(x):
    return x.data_ptr() if _is_torch(x) else x.ctypes.data


def _stream(x):
    import torch
    return torch.cuda.current_stream(x.device).cuda_stream


def _on_device(img: _Img):
    """Context making the tensor's device current for the C call."""
    import torch
    return torch.cuda.device(img.a.device)


# ---- the operations ---------------------------------------------------------------
def resize(image, nwidth: int, nheight: int, filter: FilterType, *, acc_init: float = ACC_INIT_REFERENCE,
           flags: int = L.FP_FAST):
    """imageops::resize.  `acc_init=1.0` is the reference's literal arithmetic (its vertical
    accumulator starts at 1.0); pass 0.0 for the mathematically intended filter."""
    im = _Img(image)
    out = im.empty_like(h=nheight, w=nwidth)
    lib = L.lib()
    if im.torch:
        with _on_device(im):
            _call(lib.b200img_resize_dev, im.ptr, im.w, im.h, im.c, im.dt, _ptr(out), nwidth, nheight,
                  int(filter), float(acc_init), flags, _stream(im.a))
    else:
        _call(lib.b200img_resize, im.ptr, im.w, im.h, im.c, im.dt, _ptr(out), nwidth, nheight,
              int(filter), float(acc_init), flags)
    return out


def resize_batch(images, nwidth: int, nheight: int, filter: FilterType, *, acc_init: float = ACC_INIT_REFERENCE,
                 flags: int = L.FP_FAST, out=None):
    """Device-resident batch variant: `images` is a CUDA tensor (n, h, w, c); one launch."""
    import torch
    if not (_is_torch(images) and images.is_cuda and images.dim() == 4):
        raise TypeError("resize_batch takes a CUDA tensor of shape (n, h, w, c)")
    images = images.contiguous()
    n, h, w, c = images.shape
    dt = {torch.uint8: L.U8, torch.uint16: L.U16, torch.float32: L.F32}[images.dtype]
    if out is None:
        out = torch.empty((n, nheight, nwidth, c), dtype=images.dtype, device=images.device)
    with torch.cuda.device(images.device):
        _call(L.lib().b200img_resize_batch_dev, images.data_ptr(), 0, w, h, c, dt, out.data_ptr(), 0,
              nwidth, nheight, n, int(filter), float(acc_init), flags, _stream(images))
    return out


def _gauss(image, params: GaussianBlurParameters, path: int, flags: int):
    im = _Img(image)
    out = im.empty_like()
    lib = L.lib()
    if im.torch:
        with _on_device(im):
            _call(lib.b200img_gaussian_blur_dev, im.ptr, im.w, im.h, im.c, im.dt, _ptr(out), C.byref(params.raw),
                  path, flags, _stream(im.a))
    else:
        _call(lib.b200img_gaussian_blur, im.ptr, im.w, im.h, im.c, im.dt, _ptr(out), C.byref(params.raw), path, flags)
    return out


def blur(image, radius: float, *, flags: int = L.FP_FAST):
    """imageops::blur -- NOTE `radius`, not sigma, at this revision of the reference
    (sample.rs:1034-1039): kernel = 2r+1 taps, sigma = 0.3*((k-1)/2-1)+0.8."""
    return _gauss(image, GaussianBlurParameters.new_from_radius(radius), L.BLUR_GENERIC, flags)


def blur_advanced(image, parameters: GaussianBlurParameters, *, flags: int = L.FP_FAST):
    """imageops::blur_advanced (f32 engine for every pixel type)."""
    return _gauss(image, parameters, L.BLUR_GENERIC, flags)


def blur_dyn(image, parameters: GaussianBlurParameters, *, flags: int = L.FP_FAST):
    """The arithmetic of DynamicImage::blur / blur_advanced (Q0.15 for u8/u16; sample.rs:1297)."""
    return _gauss(image, parameters, L.BLUR_DYN, flags)


def fast_blur(image, sigma: float):
    """imageops::fast_blur."""
    im = _Img(image)
    out = im.empty_like()
    lib = L.lib()
    if im.torch:
        with _on_device(im):
            _call(lib.b200img_fast_blur_dev, im.ptr, im.w, im.h, im.c, im.dt, _ptr(out), float(sigma), _stream(im.a))
    else:
        _call(lib.b200img_fast_blur, im.ptr, im.w, im.h, im.c, im.dt, _ptr(out), float(sigma))
    if out.size == 0:
        return image.clone() if im.torch else np.array(image, copy=True)
    return out


def thumbnail(image, new_width: int, new_height: int, *, flags: int = 0):
    """imageops::thumbnail (src/imageops/sample.rs:595).  A CUDA tensor of shape (n, h, w, c) is a batch."""
    if _is_torch(image) and image.dim() == 4:
        import torch
        image = image.contiguous()
        n, h, w, c = image.shape
        dt = {torch.uint8: L.U8, torch.uint16: L.U16, torch.float32: L.F32}[image.dtype]
        out = torch.zeros((n, new_height, new_width, c), dtype=image.dtype, device=image.device)
        with torch.cuda.device(image.device):
            _call(L.lib().b200img_thumbnail_batch_dev, image.data_ptr(), w, h, c, dt, out.data_ptr(), new_width, new_height,
                  n, flags, _stream(image))
        return out
    im = _Img(image)
    out = im.empty_like(h=new_height, w=new_width)
    lib = L.lib()
    if im.torch:
        with _on_device(im):
            _call(lib.b200img_thumbnail_dev, im.ptr, im.w, im.h, im.c, im.dt, _ptr(out), new_width, new_height, flags,
                  _stream(im.a))
    else:
        _call(lib.b200img_thumbnail, im.ptr, im.w, im.h, im.c, im.dt, _ptr(out), new_width, new_height, flags)
    return out


def filter3x3(image, kernel):
    """imageops::filter3x3 (src/imageops/sample.rs:897): `kernel` = 9 weights, row-major."""
    im = _Img(image)
    out = im.empty_like()
    k = (C.c_float * 9)(*[float(v) for v in np.asarray(kernel, np.float32).reshape(9)])
    lib = L.lib()
    if im.torch:
        with _on_device(im):
            _call(lib.b200img_filter3x3_dev, im.ptr, im.w, im.h, im.c, im.dt, _ptr(out), C.cast(k, C.c_void_p), _stream(im.a))
    else:
        _call(lib.b200img_filter3x3, im.ptr, im.w, im.h, im.c, im.dt, _ptr(out), C.cast(k, C.c_void_p))
    return out


def unsharpen(image, sigma: float, threshold: int, *, flags: int = L.FP_FAST):
    """imageops::unsharpen (src/imageops/sample.rs:1617)."""
    im = _Img(image)
    out = im.empty_like()
    lib = L.lib()
    if im.torch:
        with _on_device(im):
            _call(lib.b200img_unsharpen_dev, im.ptr, im.w, im.h, im.c, im.dt, _ptr(out), float(sigma), int(threshold), flags,
                  _stream(im.a))
    else:
        _call(lib.b200img_unsharpen, im.ptr, im.w, im.h, im.c, im.dt, _ptr(out), float(sigma), int(threshold), flags)
    return out


def has_constant_alpha(image) -> bool:
    """has_constant_alpha (src/imageops/resize.rs:155-218)."""
    im = _Img(image)
    res = C.c_int(1)
    lib = L.lib()
    if im.torch:
        with _on_device(im):
            _call(lib.b200img_has_constant_alpha_dev, im.ptr, im.w, im.h, im.c, im.dt, C.byref(res), _stream(im.a))
    else:
        _call(lib.b200img_has_constant_alpha, im.ptr, im.w, im.h, im.c, im.dt, C.byref(res))
    return bool(res.value)


def _alpha_mul(name: str, image):
    im = _Img(image, inplace=True)
    lib = L.lib()
    if im.torch:
        with _on_device(im):
            _call(getattr(lib, f"b200img_{name}_dev"), im.ptr, im.w, im.h, im.c, im.dt, _stream(im.a))
    else:
        _call(getattr(lib, f"b200img_{name}"), im.ptr, im.w, im.h, im.c, im.dt)
    return image


def premultiply_alpha(image):
    """In place.  The reference calls pic-scale-safe here (resize.rs:105-129): parity unpinned."""
    return _alpha_mul("premultiply_alpha", image)


def unpremultiply_alpha(image):
    """In place (resize.rs:132-151); parity unpinned."""
    return _alpha_mul("unpremultiply_alpha", image)


def convert_resize(image, dst_channels: int, dst_dtype, nwidth: int, nheight: int, filter: FilterType, *,
                   acc_init: float = ACC_INIT_REFERENCE, flags: int = L.FP_FAST):
    """Decode hand-off for a HOST image: upload once, DynamicImage::to_* on the device, resize, download."""
    a = np.ascontiguousarray(image)
    if a.ndim == 2:
        a = a[:, :, None]
    h, w, c = a.shape
    out = np.empty((nheight, nwidth, dst_channels), np.dtype(dst_dtype))
    _call(L.lib().b200img_convert_resize, a.ctypes.data, _NP2DT[a.dtype], c, w, h, out.ctypes.data, _NP2DT[out.dtype],
          dst_channels, nwidth, nheight, int(filter), float(acc_init), flags)
    return out


def _colorop(name: str, image, *params, out=None):
    im = _Img(image, inplace=out is image)
    if out is None:
        out = im.empty_like()
    lib = L.lib()
    if im.torch:
        with _on_device(im):
            _call(getattr(lib, f"b200img_{name}_dev"), im.ptr, im.w, im.h, im.c, im.dt, _ptr(out), *params, _stream(im.a))
    else:
        _call(getattr(lib, f"b200img_{name}"), im.ptr, im.w, im.h, im.c, im.dt, _ptr(out), *params)
    return out


def invert(image):
    """imageops::invert (src/imageops/colorops.rs:70) -- IN PLACE, like the reference."""
    return _colorop("invert", image, out=image)


def brighten(image, value: int):
    """imageops::brighten (colorops.rs:161)."""
    return _colorop("brighten", image, int(value))


def brighten_in_place(image, value: int):
    return _colorop("brighten", image, int(value), out=image)


def contrast(image, contrast: float):
    """imageops::contrast (colorops.rs:92)."""
    return _colorop("contrast", image, float(contrast))


def contrast_in_place(image, contrast: float):
    return _colorop("contrast", image, float(contrast), out=image)


def huerotate(image, value: int):
    """imageops::huerotate (colorops.rs:232); `value` in degrees."""
    return _colorop("huerotate", image, int(value))


def huerotate_in_place(image, value: int):
    return _colorop("huerotate", image, int(value), out=image)


def grayscale(image, *, alpha: bool = False):
    """imageops::grayscale / grayscale_alpha (colorops.rs:14-25)."""
    im = _Img(image)
    out = im.empty_like(c=2 if alpha else 1)
    lib = L.lib()
    if im.torch:
        with _on_device(im):
            _call(lib.b200img_grayscale_dev, im.ptr, im.w, im.h, im.c, im.dt, _ptr(out), int(alpha), _stream(im.a))
    else:
        _call(lib.b200img_grayscale, im.ptr, im.w, im.h, im.c, im.dt, _ptr(out), int(alpha))
    return out


def grayscale_alpha(image):
    return grayscale(image, alpha=True)


def convert_pixels(image, dst_channels: int, dst_dtype, *, mode: int = L.CAST, coefs=None):
    """Pixel-format conversion of a (h, w, c) / (n, c) array: mode CAST = DynamicImage::to_*,
    mode CONVERT = ImageBuffer::convert."""
    torch_in = _is_torch(image)
    if torch_in:
        import torch
        a = image.contiguous()
        sdt = {torch.uint8: L.U8, torch.uint16: L.U16, torch.float32: L.F32}[a.dtype]
        if isinstance(dst_dtype, torch.dtype):
            tdt = dst_dtype
        else:
            tdt = {np.dtype(np.uint8): torch.uint8, np.dtype(np.uint16): torch.uint16,
                   np.dtype(np.float32): torch.float32}[np.dtype(dst_dtype)]
        ddt = {torch.uint8: L.U8, torch.uint16: L.U16, torch.float32: L.F32}[tdt]
    else:
        a = np.ascontiguousarray(image)
        sdt = _NP2DT[a.dtype]
        ddt = _NP2DT[np.dtype(dst_dtype)]
    sc = a.shape[-1]
    npx = int(np.prod(a.shape[:-1])) if a.ndim > 1 else 0
    oshape = tuple(a.shape[:-1]) + (dst_channels,)
    cf = None
    if coefs is not None:
        cf = (C.c_float * 3)(*[float(x) for x in coefs])
    lib = L.lib()
    if torch_in:
        out = torch.empty(oshape, dtype=tdt, device=a.device)
        with torch.cuda.device(a.device):
            _call(lib.b200img_convert_dev, a.data_ptr(), sdt, sc, out.data_ptr(), ddt, dst_channels, npx, mode,
                  C.cast(cf, C.c_void_p) if cf is not None else None, _stream(a))
    else:
        out = np.empty(oshape, dtype=np.dtype(dst_dtype))
        _call(lib.b200img_convert, a.ctypes.data, sdt, sc, out.ctypes.data, ddt, dst_channels, npx, mode,
              C.cast(cf, C.c_void_p) if cf is not None else None)
    return out


def swizzle_02(buf, channels: int):
    """In place: BgraSwizzle (src/primitive_sealed.rs:34-67) on a flat subpixel buffer."""
    lib = L.lib()
    if _is_torch(buf):
        import torch
        if not buf.is_contiguous():
            raise ValueError("swizzle_02 works in place and needs a contiguous tensor")
        dt = {torch.uint8: L.U8, torch.uint16: L.U16, torch.float32: L.F32}[buf.dtype]
        with torch.cuda.device(buf.device):
            _call(lib.b200img_swizzle_02_dev, buf.data_ptr(), dt, channels, buf.numel(), _stream(buf))
    else:
        if not (isinstance(buf, np.ndarray) and buf.flags.c_contiguous and buf.flags.writeable):
            raise ValueError("swizzle_02 works in place and needs a writable contiguous ndarray")
        _call(lib.b200img_swizzle_02, buf.ctypes.data, _NP2DT[buf.dtype], channels, buf.size)
    return buf


# ---- host tables (exposed for tests and for callers that want the plan) -----------------
def build_weights(filter: FilterType, src_len: int, dst_len: int, axis: int):
    lib = L.lib()
    mt = max(lib.b200img_resize_max_taps(int(filter), src_len, dst_len), 1)
    left = np.zeros(dst_len, np.int32)
    cnt = np.zeros(dst_len, np.int32)
    w = np.zeros((dst_len, mt), np.float32)
    _call(lib.b200img_build_weights, int(filter), src_len, dst_len, axis, left.ctypes.data, cnt.ctypes.data,
          w.ctypes.data, mt)
    return left, cnt, w


def gaussian_kernel_1d(width: int, sigma: float):
    out = np.zeros(width, np.float32)
    _call(L.lib().b200img_gaussian_kernel_1d, width, float(sigma), out.ctypes.data)
    return out


def boxes_for_gauss(sigma: float):
    out = (C.c_uint32 * 3)()
    _call(L.lib().b200img_boxes_for_gauss, float(sigma), out)
    return list(out)


def resize_dimensions(w: int, h: int, nw: int, nh: int, fill: bool):
    ow, oh = C.c_uint32(), C.c_uint32()
    _call(L.lib().b200img_resize_dimensions, w, h, nw, nh, int(fill), C.byref(ow), C.byref(oh))
    return ow.value, oh.value
