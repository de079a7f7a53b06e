"""CPU-side checks of the product: the C-ABI library loads and exports every symbol the
header declares, and its host-side tables equal the oracle's bit for bit.  No GPU compute."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import oracle as o
from image_b200 import FilterType, GaussianBlurParameters, PanicError, imageops
from image_b200 import _lib as L

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_header_symbols_exported():
    hdr = open(os.path.join(ROOT, "include", "b200img.h")).read()
    declared = set(re.findall(r"\b(b200img_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 30
    lib = L.lib()
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in b200img.h but not exported"
    assert declared == set(L.SIGNATURES), declared ^ set(L.SIGNATURES)
    assert lib.b200img_version() == 100


@pytest.mark.parametrize("f", list(FilterType))
@pytest.mark.parametrize("src,dst", [(1080, 540), (1920, 960), (1024, 256), (100, 37), (16, 64), (629, 22),
                                     (7, 7), (1, 5), (5, 1), (32768, 65536 // 64), (3, 1000)])
def test_weight_tables_equal_oracle(f, src, dst):
    for axis in (0, 1):
        a = imageops.build_weights(f, src, dst, axis)
        b = o.resize_weights(int(f), src, dst, axis)
        n = min(a[2].shape[1], b[2].shape[1])
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
        assert np.array_equal(a[2][:, :n], b[2][:, :n])
        assert (a[1] >= 1).all() and (a[0] >= 0).all() and (a[0] + a[1] <= src).all()


def test_gauss_tables_equal_oracle():
    for r in (0.0, 0.3, 1.0, 2.5, 10.0, 33.3):
        a, b = GaussianBlurParameters.new_from_radius(r).raw, o.gauss_params_from_radius(r)
        assert (a.x_kernel_size, a.x_sigma, a.y_kernel_size, a.y_sigma) == (b.x_kernel_size, b.x_sigma, b.y_kernel_size, b.y_sigma)
    for s in (0.0, 0.1, 0.17, 0.5, 3.0, 7.0, 10.0, 50.0):
        a, b = GaussianBlurParameters.new_from_sigma(s).raw, o.gauss_params_from_sigma(s)
        assert (a.x_kernel_size, a.x_sigma) == (b.x_kernel_size, b.x_sigma)
        assert np.array_equal(imageops.gaussian_kernel_1d(a.x_kernel_size, a.x_sigma),
                              o.gaussian_kernel_1d(b.x_kernel_size, b.x_sigma))
    a, b = GaussianBlurParameters.new_anisotropic_kernel_size(4.2, 9.9).raw, o.gauss_params_from_anisotropic(4.2, 9.9)
    assert (a.x_kernel_size, a.x_sigma, a.y_kernel_size, a.y_sigma) == (b.x_kernel_size, b.x_sigma, b.y_kernel_size, b.y_sigma)
    for s in (0.0, 0.5, 1.0, 3.3, 10.0, 50.0, 99.0):
        assert imageops.boxes_for_gauss(s) == o.boxes_for_gauss(s)


def test_parameter_panics_become_errors():
    # asserts in src/imageops/sample.rs:1134-1158, 1211-1214
    for bad in (-1.0, float("nan"), float("inf")):
        with pytest.raises(PanicError):
            GaussianBlurParameters.new_from_radius(bad)
        with pytest.raises(PanicError):
            GaussianBlurParameters.new_from_sigma(bad)
    with pytest.raises(PanicError):
        GaussianBlurParameters.new_from_kernel_size(0.0)
    assert "gauss_params" in L.last_error()


def test_resize_dimensions_equal_oracle():
    rng = np.random.default_rng(4)
    for _ in range(200):
        w, h, nw, nh = (int(v) for v in rng.integers(1, 5000, 4))
        for fill in (False, True):
            assert imageops.resize_dimensions(w, h, nw, nh, fill) == o.resize_dimensions(w, h, nw, nh, fill)
    assert imageops.resize_dimensions(1, 1, 2 ** 32 - 1, 7, True) == o.resize_dimensions(1, 1, 2 ** 32 - 1, 7, True)


def test_argument_errors_without_gpu():
    lib = L.lib()
    buf = (C.c_uint8 * 64)()
    assert lib.b200img_resize(buf, 4, 4, 5, L.U8, buf, 2, 2, 0, 1.0, 0) == L.ERR_BAD_ARG       # 5 channels
    assert lib.b200img_resize(buf, 4, 4, 3, 9, buf, 2, 2, 0, 1.0, 0) == L.ERR_BAD_ARG          # dtype
    assert lib.b200img_resize(buf, 4, 4, 3, L.U8, buf, 2, 2, 17, 1.0, 0) == L.ERR_BAD_ARG      # filter
    assert lib.b200img_resize(buf, 2 ** 32 - 1, 2 ** 32 - 1, 4, L.F32, buf, 2, 2, 0, 1.0, 0) == L.ERR_DIMENSION
    p = L.GaussParams(4, 1.0, 3, 1.0)  # even kernel: assert in filter_1d.rs:649-654
    assert lib.b200img_gaussian_blur(buf, 2, 2, 3, L.U8, buf, C.byref(p), 0, 0) == L.ERR_BAD_ARG
    assert lib.b200img_swizzle_02(buf, L.U8, 2, 64) == L.ERR_BAD_ARG
    assert lib.b200img_convert(buf, L.U8, 0, buf, L.U8, 3, 4, 0, None) == L.ERR_BAD_ARG
    # empty work is fine without touching a device
    assert lib.b200img_resize(buf, 4, 4, 3, L.U8, buf, 0, 0, 0, 1.0, 0) == 0
    assert lib.b200img_fast_blur(buf, 0, 0, 3, L.U8, buf, 2.0) == 0
    assert lib.b200img_convert(buf, L.U8, 3, buf, L.U8, 4, 0, 0, None) == 0


def test_no_cpu_fallback():
    """Without a GPU every compute entry point fails loudly -- it never computes on the host."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    img = np.zeros((8, 8, 3), np.uint8)
    with pytest.raises(L.B200ImgError) as e:
        imageops.resize(img, 4, 4, FilterType.Triangle)
    assert e.value.status in (L.ERR_CUDA, L.ERR_OOM)
    with pytest.raises(L.B200ImgError):
        imageops.fast_blur(img, 2.0)
    with pytest.raises(L.B200ImgError):
        imageops.convert_pixels(img, 4, np.uint16)


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "image_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                txt = open(os.path.join(dp, f), errors="replace").read()
                assert "import oracle" not in txt and "liboracle" not in txt and "oracle/" not in txt, f


def test_rust_ffi_declares_every_header_symbol():
    """rust/image-b200/src/ffi.rs (authored, not compilable here) stays in step with the header."""
    hdr = open(os.path.join(ROOT, "include", "b200img.h")).read()
    rs = open(os.path.join(ROOT, "rust", "image-b200", "src", "ffi.rs")).read()
    declared = set(re.findall(r"\b(b200img_[a-z0-9_]+)\s*\(", hdr))
    bound = set(re.findall(r"fn (b200img_[a-z0-9_]+)", rs))
    assert declared == bound, declared ^ bound


def test_thumbnail_reference_panic_is_refused_without_a_gpu():
    """thumbnail_sample_block divides by (w/256)*(h/256) = 0 (sample.rs:716,739): BAD_ARG before any CUDA call"""
    lib = L.lib()
    src = np.zeros((1000, 100, 1), np.uint8)
    dst = np.zeros((1, 1, 1), np.uint8)
    rc = lib.b200img_thumbnail(src.ctypes.data, 100, 1000, 1, L.U8, dst.ctypes.data, 1, 1, 0)
    assert rc == L.ERR_BAD_ARG and "divides by zero" in L.last_error()
    assert lib.b200img_thumbnail(src.ctypes.data, 100, 1000, 5, L.U8, dst.ctypes.data, 1, 1, 0) == L.ERR_BAD_ARG
    # an empty destination is a no-op
    assert lib.b200img_thumbnail(src.ctypes.data, 100, 1000, 1, L.U8, dst.ctypes.data, 0, 7, 0) == 0
