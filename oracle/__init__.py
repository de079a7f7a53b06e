"""CPU oracle for the image-rs pixel hot path -- TEST INFRASTRUCTURE ONLY.

ctypes bindings over ``oracle/_build/liboracle.so`` (C++ restatement of the
reference's Rust, see ``oracle.h`` for citations).  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline legs may import
this package; ``image_b200`` never does.

Images are numpy arrays of shape (h, w, c) (or (h, w) for one channel) and
dtype uint8 / uint16 / float32, C-contiguous.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "liboracle.so")

NEAREST, TRIANGLE, CATMULLROM, GAUSSIAN, LANCZOS3 = range(5)
FILTERS = {"nearest": NEAREST, "triangle": TRIANGLE, "catmullrom": CATMULLROM,
           "gaussian": GAUSSIAN, "lanczos3": LANCZOS3}
CAST, CONVERT = 0, 1
_DT = {np.dtype(np.uint8): 0, np.dtype(np.uint16): 1, np.dtype(np.float32): 2}
_NP = {0: np.uint8, 1: np.uint16, 2: np.float32}


def build(force: bool = False) -> str:
    """Compile the oracle with its Makefile (gcc only; no GPU, no reference sources)."""
    srcs = [os.path.join(_HERE, f) for f in os.listdir(_HERE) if f.endswith((".cpp", ".h")) or f == "Makefile"]
    stale = (not os.path.exists(_SO)) or any(os.path.getmtime(s) > os.path.getmtime(_SO) for s in srcs)
    if force or stale:
        subprocess.check_call(["make", "-C", _HERE] + (["-B"] if force else []),
                              stdout=subprocess.DEVNULL)
    return _SO


class GaussParams(C.Structure):
    _fields_ = [("x_kernel_size", C.c_uint32), ("x_sigma", C.c_float),
                ("y_kernel_size", C.c_uint32), ("y_sigma", C.c_float)]

    def __repr__(self):
        return (f"GaussParams(x=({self.x_kernel_size}, {self.x_sigma}), "
                f"y=({self.y_kernel_size}, {self.y_sigma}))")


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_SO)
        vp, u32, i, f, sz = C.c_void_p, C.c_uint32, C.c_int, C.c_float, C.c_size_t
        L.orc_set_fused_mac.argtypes = [i]
        L.orc_filter_support.argtypes = [i]; L.orc_filter_support.restype = f
        L.orc_filter_kernel.argtypes = [i, f]; L.orc_filter_kernel.restype = f
        L.orc_resize_max_taps.argtypes = [i, u32, u32]
        L.orc_resize_weights.argtypes = [i, u32, u32, i, vp, vp, vp, i]
        L.orc_resize.argtypes = [vp, u32, u32, i, i, vp, u32, u32, i, f]
        L.orc_resize_mt.argtypes = [vp, u32, u32, i, i, vp, u32, u32, i, f, i]
        P = C.POINTER(GaussParams)
        L.orc_gauss_params_from_radius.argtypes = [f, P]
        L.orc_gauss_params_from_kernel_size.argtypes = [f, P]
        L.orc_gauss_params_from_anisotropic.argtypes = [f, f, P]
        L.orc_gauss_params_from_sigma.argtypes = [f, P]
        L.orc_gaussian_kernel_1d.argtypes = [u32, f, vp]; L.orc_gaussian_kernel_1d.restype = None
        L.orc_gaussian_blur_generic.argtypes = [vp, u32, u32, i, i, vp, P]
        L.orc_gaussian_blur_dyn.argtypes = [vp, u32, u32, i, i, vp, P]
        L.orc_gaussian_blur_generic_chunked.argtypes = [vp, u32, u32, i, i, vp, P]
        L.orc_gaussian_blur_dyn_chunked.argtypes = [vp, u32, u32, i, i, vp, P]
        L.orc_q15_kernel.argtypes = [vp, i, vp]
        L.orc_boxes_for_gauss.argtypes = [f, vp]; L.orc_boxes_for_gauss.restype = None
        L.orc_fast_blur.argtypes = [vp, u32, u32, i, i, vp, f]
        L.orc_u8_box_store.argtypes = [u32, u32]; L.orc_u8_box_store.restype = C.c_uint8
        L.orc_convert.argtypes = [vp, i, i, vp, i, i, sz, i, vp]
        L.orc_swizzle_02.argtypes = [vp, i, i, sz]
        L.orc_filter3x3.argtypes = [vp, u32, u32, i, i, vp, vp]
        L.orc_unsharpen.argtypes = [vp, u32, u32, i, i, vp, f, C.c_int32]
        L.orc_thumbnail.argtypes = [vp, u32, u32, i, i, vp, u32, u32]
        L.orc_has_constant_alpha.argtypes = [vp, u32, u32, i, i]
        L.orc_alpha_mul.argtypes = [vp, sz, i, i, i]
        L.orc_colorop.argtypes = [i, vp, vp, sz, i, i, C.c_int32, f]
        L.orc_huerotate_matrix.argtypes = [C.c_int32, vp]; L.orc_huerotate_matrix.restype = None
        pu32 = C.POINTER(u32)
        L.orc_resize_dimensions.argtypes = [u32, u32, u32, u32, i, pu32, pu32]
        L.orc_resize_dimensions.restype = None
        _lib = L
    return _lib


def _img(a):
    a = np.ascontiguousarray(a)
    if a.dtype not in _DT:
        raise TypeError(f"unsupported dtype {a.dtype}")
    if a.ndim == 2:
        a = a[:, :, None]
    if a.ndim != 3 or not 1 <= a.shape[2] <= 4:
        raise ValueError(f"expected (h, w, c<=4), got {a.shape}")
    return a


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def _filter(f):
    return FILTERS[f.lower()] if isinstance(f, str) else int(f)


def set_fused_mac(fused: bool):
    lib().orc_set_fused_mac(int(bool(fused)))


def filter_support(f) -> float:
    return lib().orc_filter_support(_filter(f))


def filter_kernel(f, x: float) -> float:
    return lib().orc_filter_kernel(_filter(f), float(x))


def resize_weights(f, src_len: int, dst_len: int, axis: int):
    """(left[int32], count[int32], weights[dst_len, max_taps] f32); axis 0 = vertical rule, 1 = horizontal."""
    L = lib()
    mt = max(L.orc_resize_max_taps(_filter(f), src_len, dst_len), 1)
    left = np.zeros(dst_len, np.int32)
    cnt = np.zeros(dst_len, np.int32)
    w = np.zeros((dst_len, mt), np.float32)
    rc = L.orc_resize_weights(_filter(f), src_len, dst_len, axis, _ptr(left), _ptr(cnt), _ptr(w), mt)
    if rc:
        raise RuntimeError(f"orc_resize_weights rc={rc}")
    return left, cnt, w


def resize(img, nw: int, nh: int, f, acc_init: float = 1.0, threads: int = 1):
    """imageops::resize (src/imageops/sample.rs:970).  acc_init=1.0 is the literal reference."""
    a = _img(img)
    h, w, c = a.shape
    out = np.zeros((nh, nw, c), a.dtype)
    rc = lib().orc_resize_mt(_ptr(a), w, h, c, _DT[a.dtype], _ptr(out), nw, nh, _filter(f),
                             float(acc_init), int(threads))
    if rc:
        raise RuntimeError(f"orc_resize rc={rc}")
    return out


def _params(fn, *args) -> GaussParams:
    p = GaussParams()
    if fn(*args, C.byref(p)):
        raise ValueError("invalid Gaussian blur parameter (the reference panics here)")
    return p


def gauss_params_from_radius(r):
    return _params(lib().orc_gauss_params_from_radius, float(r))


def gauss_params_from_kernel_size(k):
    return _params(lib().orc_gauss_params_from_kernel_size, float(k))


def gauss_params_from_anisotropic(xk, yk):
    return _params(lib().orc_gauss_params_from_anisotropic, float(xk), float(yk))


def gauss_params_from_sigma(s):
    return _params(lib().orc_gauss_params_from_sigma, float(s))


def gaussian_kernel_1d(width: int, sigma: float):
    out = np.zeros(width, np.float32)
    lib().orc_gaussian_kernel_1d(width, float(sigma), _ptr(out))
    return out


def q15_kernel(k):
    k = np.ascontiguousarray(k, np.float32)
    out = np.zeros(len(k), np.uint32)
    n = lib().orc_q15_kernel(_ptr(k), len(k), _ptr(out))
    return out[:n]


def gaussian_blur(img, params: GaussParams, path: str = "generic", chunked: bool = False):
    """path 'generic' = imageops::blur/blur_advanced (f32 engine);
    'dyn' = DynamicImage::blur/blur_advanced (Q0.15 for u8/u16).
    chunked = the filters with the reference's chunk structure (vectorised by the compiler, as rustc
    vectorises the reference's): bit-identical, several times faster -- what the CPU baseline times."""
    a = _img(img)
    h, w, c = a.shape
    out = np.zeros_like(a)
    L = lib()
    if chunked:
        fn = L.orc_gaussian_blur_generic_chunked if path == "generic" else L.orc_gaussian_blur_dyn_chunked
    else:
        fn = L.orc_gaussian_blur_generic if path == "generic" else L.orc_gaussian_blur_dyn
    rc = fn(_ptr(a), w, h, c, _DT[a.dtype], _ptr(out), C.byref(params))
    if rc:
        raise RuntimeError(f"gaussian blur rc={rc}")
    return out


def boxes_for_gauss(sigma: float):
    out = (C.c_size_t * 3)()
    lib().orc_boxes_for_gauss(float(sigma), out)
    return list(out)


def fast_blur(img, sigma: float):
    a = _img(img)
    h, w, c = a.shape
    out = a.copy()
    rc = lib().orc_fast_blur(_ptr(a), w, h, c, _DT[a.dtype], _ptr(out), float(sigma))
    if rc:
        raise RuntimeError(f"orc_fast_blur rc={rc}")
    return out


def u8_box_store(acc: int, ks: int) -> int:
    return lib().orc_u8_box_store(acc, ks)


def convert(pixels, src_channels: int, dst_dtype, dst_channels: int, mode: int = CAST, coefs=None):
    """pixels: flat or (n, src_channels) array.  Returns (n, dst_channels)."""
    a = np.ascontiguousarray(pixels).reshape(-1, src_channels)
    n = a.shape[0]
    out = np.zeros((n, dst_channels), np.dtype(dst_dtype))
    cf = None
    if coefs is not None:
        cf = np.ascontiguousarray(coefs, np.float32)
    rc = lib().orc_convert(_ptr(a), _DT[a.dtype], src_channels, _ptr(out), _DT[out.dtype],
                           dst_channels, n, mode, _ptr(cf) if cf is not None else None)
    if rc:
        raise RuntimeError(f"orc_convert rc={rc}")
    return out


def swizzle_02(buf, channels: int):
    """In place on a copy: returns the swizzled flat array (trailing partial chunk untouched)."""
    a = np.ascontiguousarray(buf).copy().reshape(-1)
    rc = lib().orc_swizzle_02(_ptr(a), _DT[a.dtype], channels, a.size)
    if rc:
        raise RuntimeError(f"orc_swizzle_02 rc={rc}")
    return a


def resize_dimensions(w, h, nw, nh, fill: bool):
    ow, oh = C.c_uint32(), C.c_uint32()
    lib().orc_resize_dimensions(w, h, nw, nh, int(fill), C.byref(ow), C.byref(oh))
    return ow.value, oh.value


def filter3x3(img, kernel):
    """imageops::filter3x3 (src/imageops/sample.rs:897)."""
    a = _img(img)
    h, w, c = a.shape
    k = np.ascontiguousarray(kernel, np.float32).reshape(9)
    out = np.zeros_like(a)
    rc = lib().orc_filter3x3(_ptr(a), w, h, c, _DT[a.dtype], _ptr(out), _ptr(k))
    if rc:
        raise RuntimeError(f"orc_filter3x3 rc={rc}")
    return out


def unsharpen(img, sigma: float, threshold: int):
    """imageops::unsharpen (src/imageops/sample.rs:1617)."""
    a = _img(img)
    h, w, c = a.shape
    out = np.zeros_like(a)
    rc = lib().orc_unsharpen(_ptr(a), w, h, c, _DT[a.dtype], _ptr(out), float(sigma), int(threshold))
    if rc:
        raise RuntimeError(f"orc_unsharpen rc={rc}")
    return out


class ReferencePanic(RuntimeError):
    """The reference panics on this input (integer division by zero in thumbnail_sample_block)."""


def thumbnail(img, nw: int, nh: int):
    """imageops::thumbnail (src/imageops/sample.rs:595)."""
    a = _img(img)
    h, w, c = a.shape
    out = np.zeros((nh, nw, c), a.dtype)
    rc = lib().orc_thumbnail(_ptr(a), w, h, c, _DT[a.dtype], _ptr(out), nw, nh)
    if rc == -2:
        raise ReferencePanic("thumbnail: block of > 65536 pixels under 256 wide or high")
    if rc:
        raise RuntimeError(f"orc_thumbnail rc={rc}")
    return out


def _colorop(op, img, ivalue=0, fvalue=0.0):
    a = _img(img)
    h, w, c = a.shape
    out = np.zeros_like(a)
    rc = lib().orc_colorop(op, _ptr(a), _ptr(out), h * w, c, _DT[a.dtype], int(ivalue), float(fvalue))
    if rc:
        raise RuntimeError(f"orc_colorop rc={rc}")
    return out


def invert(img):
    """imageops::invert (src/imageops/colorops.rs:70)."""
    return _colorop(0, img)


def brighten(img, value: int):
    """imageops::brighten (colorops.rs:161)."""
    return _colorop(1, img, ivalue=value)


def contrast(img, contrast: float):
    """imageops::contrast (colorops.rs:92)."""
    return _colorop(2, img, fvalue=contrast)


def huerotate(img, degrees: int):
    """imageops::huerotate (colorops.rs:232)."""
    return _colorop(3, img, ivalue=degrees)


def huerotate_matrix(degrees: int):
    m = np.zeros(9, np.float64)
    lib().orc_huerotate_matrix(int(degrees), _ptr(m))
    return m


def grayscale(img, alpha: bool = False):
    """imageops::grayscale / grayscale_alpha (colorops.rs:14-66): pixel.to_luma() in the source type."""
    a = _img(img)
    h, w, c = a.shape
    return convert(a.reshape(-1, c), c, a.dtype, 2 if alpha else 1, CONVERT).reshape(h, w, 2 if alpha else 1)


def has_constant_alpha(img) -> bool:
    """has_constant_alpha (src/imageops/resize.rs:155-218)."""
    a = _img(img)
    h, w, c = a.shape
    return bool(lib().orc_has_constant_alpha(_ptr(a), w, h, c, _DT[a.dtype]))


def alpha_mul(img, un: bool):
    """The product's own premultiply / unpremultiply rule (the reference's is pic-scale-safe's: unpinned)."""
    a = _img(img).copy()
    h, w, c = a.shape
    lib().orc_alpha_mul(_ptr(a), h * w, c, _DT[a.dtype], int(un))
    return a


def resize_region(fetch_rows, w: int, h: int, c: int, dtype, nw: int, nh: int, f, y0: int, y1: int, x0: int, x1: int,
                  acc_init: float = 1.0):
    """Destination window [y0, y1) x [x0, x1) of imageops::resize of a w x h image that is too large to run whole:
    the C oracle's own weight tables (orc_resize_weights) applied by numpy in the reference's order -- vertical
    pass first with the accumulator starting at `acc_init`, every multiply and add rounded separately to f32, taps
    ascending (sample.rs:546-558), then the horizontal pass from 0 (310-334), then clamp_nearest_from.
    `fetch_rows(r0, r1, c0, c1)` returns source rows [r0, r1), columns [c0, c1) as an array (rows, cols, c).
    Bit-identical with resize() on images small enough for both (tests/test_oracle_kat.py)."""
    dtype = np.dtype(dtype)
    vl, vc, vw = resize_weights(f, h, nh, 0)
    hl, hc, hw = resize_weights(f, w, nw, 1)
    r0, r1 = int(vl[y0]), int(vl[y1 - 1] + vc[y1 - 1])
    c0, c1 = int(hl[x0]), int(hl[x1 - 1] + hc[x1 - 1])
    src = np.asarray(fetch_rows(r0, r1, c0, c1)).astype(np.float32)           # exact for u8 / u16
    F = np.float32
    tmp = np.empty((y1 - y0, c1 - c0, c), np.float32)
    for oy in range(y0, y1):
        acc = np.full((c1 - c0, c), F(acc_init), np.float32)
        for i in range(int(vc[oy])):
            acc = acc + src[int(vl[oy]) - r0 + i] * vw[oy, i]
        tmp[oy - y0] = acc
    out = np.empty((y1 - y0, x1 - x0, c), np.float32)
    for ox in range(x0, x1):
        acc = np.zeros((y1 - y0, c), np.float32)
        for i in range(int(hc[ox])):
            acc = acc + tmp[:, int(hl[ox]) - c0 + i] * hw[ox, i]
        out[:, ox - x0] = acc
    if dtype == np.float32:
        return np.minimum(np.maximum(out, F(0.0)), F(1.0))   # max(0).min(1); a NaN sample is not expected here
    mx = np.iinfo(dtype).max
    v = out + F(0.5)
    return np.where(np.isnan(v), 0, np.clip(np.trunc(v), 0, mx)).astype(dtype)
