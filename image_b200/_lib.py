"""ctypes binding of libb200img.so (the C ABI in include/b200img.h).

The library is the product: if it is missing or cannot be loaded this module
raises -- there is no Python or CPU fallback for any pixel operation.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.path.join(_HERE, "lib", "libb200img.so")

U8, U16, F32 = 0, 1, 2
NEAREST, TRIANGLE, CATMULLROM, GAUSSIAN, LANCZOS3 = range(5)
FP_FAST, FP_STRICT, FP_FUSED_MAC, RESIZE_UNFUSED, RESIZE_TILED, THUMBNAIL_PER_PIXEL = 0, 1, 2, 256, 512, 1024
GAUSS_TWO_KERNELS, GAUSS_FUSED = 2048, 4096
RESIZE_GENERAL = 8192
BLUR_GENERIC, BLUR_DYN = 0, 1
CAST, CONVERT = 0, 1
OK, ERR_BAD_ARG, ERR_DIMENSION, ERR_OOM, ERR_CUDA, ERR_UNSUPPORTED = 0, -1, -2, -3, -4, -5


class GaussParams(C.Structure):
    """b200img_gauss_params == GaussianBlurParameters (src/imageops/sample.rs:1081-1105)."""
    _fields_ = [("x_kernel_size", C.c_uint32), ("x_sigma", C.c_float),
                ("y_kernel_size", C.c_uint32), ("y_sigma", C.c_float)]

    def __repr__(self):
        return (f"GaussParams(x=({self.x_kernel_size}, {self.x_sigma}), "
                f"y=({self.y_kernel_size}, {self.y_sigma}))")


class Band(C.Structure):
    """b200img_band: one device's share of a row-band split (include/b200img.h)."""
    _fields_ = [("device", C.c_int), ("d_src", C.c_void_p), ("src_row0", C.c_uint32), ("src_rows", C.c_uint32),
                ("own_row0", C.c_uint32), ("own_rows", C.c_uint32), ("d_dst", C.c_void_p),
                ("dst_row0", C.c_uint32), ("dst_rows", C.c_uint32)]


class B200ImgError(RuntimeError):
    def __init__(self, status: int, message: str):
        super().__init__(f"b200img status {status}: {message}")
        self.status = status
        self.message = message


_vp, _u32, _i, _f, _sz = C.c_void_p, C.c_uint32, C.c_int, C.c_float, C.c_size_t
_P = C.POINTER(GaussParams)
_pu32 = C.POINTER(C.c_uint32)
_B = C.POINTER(Band)

# name -> (restype, argtypes); every symbol include/b200img.h declares
SIGNATURES = {
    "b200img_last_error": (C.c_char_p, []),
    "b200img_version": (_i, []),
    "b200img_device_count": (_i, [C.POINTER(_i)]),
    "b200img_pinned_alloc": (_i, [_sz, C.POINTER(_vp)]),
    "b200img_pinned_free": (_i, [_vp]),
    "b200img_release_caches": (_i, []),
    "b200img_filter_support": (_f, [_i]),
    "b200img_filter_kernel": (_f, [_i, _f]),
    "b200img_resize_max_taps": (_i, [_i, _u32, _u32]),
    "b200img_build_weights": (_i, [_i, _u32, _u32, _i, _vp, _vp, _vp, _i]),
    "b200img_resize_dimensions": (_i, [_u32, _u32, _u32, _u32, _i, _pu32, _pu32]),
    "b200img_gauss_params_from_radius": (_i, [_f, _P]),
    "b200img_gauss_params_from_kernel_size": (_i, [_f, _P]),
    "b200img_gauss_params_from_anisotropic": (_i, [_f, _f, _P]),
    "b200img_gauss_params_from_sigma": (_i, [_f, _P]),
    "b200img_gaussian_kernel_1d": (_i, [_u32, _f, _vp]),
    "b200img_boxes_for_gauss": (_i, [_f, _pu32]),
    "b200img_resize": (_i, [_vp, _u32, _u32, _i, _i, _vp, _u32, _u32, _i, _f, _i]),
    "b200img_resize_dev": (_i, [_vp, _u32, _u32, _i, _i, _vp, _u32, _u32, _i, _f, _i, _vp]),
    "b200img_resize_batch_dev": (_i, [_vp, _sz, _u32, _u32, _i, _i, _vp, _sz, _u32, _u32, _u32, _i, _f, _i, _vp]),
    "b200img_resize_batch": (_i, [_vp, _u32, _u32, _i, _i, _vp, _u32, _u32, _u32, _i, _f, _i]),
    "b200img_resize_band_src_rows": (_i, [_u32, _u32, _i, _u32, _u32, _pu32, _pu32]),
    "b200img_resize_band_dev": (_i, [_vp, _u32, _u32, _u32, _u32, _i, _i, _vp, _u32, _u32, _u32, _u32, _i, _f, _i, _vp]),
    "b200img_gaussian_blur": (_i, [_vp, _u32, _u32, _i, _i, _vp, _P, _i, _i]),
    "b200img_gaussian_blur_dev": (_i, [_vp, _u32, _u32, _i, _i, _vp, _P, _i, _i, _vp]),
    "b200img_gaussian_blur_batch": (_i, [_vp, _u32, _u32, _i, _i, _vp, _u32, _P, _i, _i]),
    "b200img_gaussian_blur_batch_dev": (_i, [_vp, _u32, _u32, _i, _i, _vp, _u32, _P, _i, _i, _vp]),
    "b200img_fast_blur": (_i, [_vp, _u32, _u32, _i, _i, _vp, _f]),
    "b200img_fast_blur_dev": (_i, [_vp, _u32, _u32, _i, _i, _vp, _f, _vp]),
    "b200img_fast_blur_batch": (_i, [_vp, _u32, _u32, _i, _i, _vp, _u32, _f]),
    "b200img_fast_blur_batch_dev": (_i, [_vp, _u32, _u32, _i, _i, _vp, _u32, _f, _vp]),
    "b200img_fast_blur_noalloc_dev": (_i, [_vp, _u32, _u32, _i, _i, _vp, _f, _vp]),
    "b200img_plan_resize_bands": (_i, [_u32, _u32, _i, _i, _B]),
    "b200img_plan_gaussian_blur_bands": (_i, [_u32, _P, _i, _B]),
    "b200img_plan_fast_blur_bands": (_i, [_u32, _f, _i, _B]),
    "b200img_exchange_ghost_rows": (_i, [_B, _i, _u32, _sz, _vp]),
    "b200img_resize_bands": (_i, [_B, _i, _u32, _u32, _i, _i, _u32, _u32, _i, _f, _i, _vp]),
    "b200img_gaussian_blur_bands": (_i, [_B, _i, _u32, _u32, _i, _i, _P, _i, _i, _vp]),
    "b200img_fast_blur_bands": (_i, [_B, _i, _u32, _u32, _i, _i, _f, _vp]),
    "b200img_gaussian_blur_band_src_rows": (_i, [_u32, _P, _u32, _u32, _pu32, _pu32]),
    "b200img_gaussian_blur_band_dev": (_i, [_vp, _u32, _u32, _u32, _u32, _i, _i, _vp, _u32, _u32, _P, _i, _i, _vp]),
    "b200img_fast_blur_band_src_rows": (_i, [_u32, _f, _u32, _u32, _pu32, _pu32]),
    "b200img_fast_blur_band_dev": (_i, [_vp, _u32, _u32, _u32, _u32, _i, _i, _vp, _u32, _u32, _f, _vp]),
    "b200img_thumbnail": (_i, [_vp, _u32, _u32, _i, _i, _vp, _u32, _u32, _i]),
    "b200img_thumbnail_dev": (_i, [_vp, _u32, _u32, _i, _i, _vp, _u32, _u32, _i, _vp]),
    "b200img_thumbnail_batch_dev": (_i, [_vp, _u32, _u32, _i, _i, _vp, _u32, _u32, _u32, _i, _vp]),
    "b200img_filter3x3": (_i, [_vp, _u32, _u32, _i, _i, _vp, _vp]),
    "b200img_filter3x3_dev": (_i, [_vp, _u32, _u32, _i, _i, _vp, _vp, _vp]),
    "b200img_unsharpen": (_i, [_vp, _u32, _u32, _i, _i, _vp, _f, _i, _i]),
    "b200img_unsharpen_dev": (_i, [_vp, _u32, _u32, _i, _i, _vp, _f, _i, _i, _vp]),
    "b200img_has_constant_alpha": (_i, [_vp, _u32, _u32, _i, _i, C.POINTER(C.c_int)]),
    "b200img_has_constant_alpha_dev": (_i, [_vp, _u32, _u32, _i, _i, C.POINTER(C.c_int), _vp]),
    "b200img_premultiply_alpha": (_i, [_vp, _u32, _u32, _i, _i]),
    "b200img_premultiply_alpha_dev": (_i, [_vp, _u32, _u32, _i, _i, _vp]),
    "b200img_unpremultiply_alpha": (_i, [_vp, _u32, _u32, _i, _i]),
    "b200img_unpremultiply_alpha_dev": (_i, [_vp, _u32, _u32, _i, _i, _vp]),
    "b200img_convert_resize": (_i, [_vp, _i, _i, _u32, _u32, _vp, _i, _i, _u32, _u32, _i, _f, _i]),
    "b200img_invert": (_i, [_vp, _u32, _u32, _i, _i, _vp]),
    "b200img_invert_dev": (_i, [_vp, _u32, _u32, _i, _i, _vp, _vp]),
    "b200img_brighten": (_i, [_vp, _u32, _u32, _i, _i, _vp, C.c_int32]),
    "b200img_brighten_dev": (_i, [_vp, _u32, _u32, _i, _i, _vp, C.c_int32, _vp]),
    "b200img_contrast": (_i, [_vp, _u32, _u32, _i, _i, _vp, _f]),
    "b200img_contrast_dev": (_i, [_vp, _u32, _u32, _i, _i, _vp, _f, _vp]),
    "b200img_huerotate": (_i, [_vp, _u32, _u32, _i, _i, _vp, C.c_int32]),
    "b200img_huerotate_dev": (_i, [_vp, _u32, _u32, _i, _i, _vp, C.c_int32, _vp]),
    "b200img_grayscale": (_i, [_vp, _u32, _u32, _i, _i, _vp, _i]),
    "b200img_grayscale_dev": (_i, [_vp, _u32, _u32, _i, _i, _vp, _i, _vp]),
    "b200img_convert": (_i, [_vp, _i, _i, _vp, _i, _i, _sz, _i, _vp]),
    "b200img_convert_dev": (_i, [_vp, _i, _i, _vp, _i, _i, _sz, _i, _vp, _vp]),
    "b200img_swizzle_02": (_i, [_vp, _i, _i, _sz]),
    "b200img_swizzle_02_dev": (_i, [_vp, _i, _i, _sz, _vp]),
}

_lib = None


def lib() -> C.CDLL:
    """The loaded library.  Raises if it has not been built (python -m image_b200.build)."""
    global _lib
    if _lib is None:
        if not os.path.exists(SO_PATH):
            raise ImportError(
                f"{SO_PATH} is missing: build it with `python -m image_b200.build` "
                "(image_b200 has no fallback path without its CUDA library)")
        L = C.CDLL(SO_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)  # AttributeError if the .so lacks a declared symbol
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def last_error() -> str:
    return lib().b200img_last_error().decode("utf-8", "replace")


def check(rc: int) -> None:
    if rc != 0:
        raise B200ImgError(rc, last_error())
