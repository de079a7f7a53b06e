"""Out-of-bounds WRITE check for every device entry point: the destination sits between two guard
bands inside a larger allocation; the bands must come back untouched and the result must equal
the one computed into a plain buffer.  (compute-sanitizer is closed on the GPU pool -- its exact answer is kept in
profiles/r02_sanitizer_closed.txt -- so this, NaN-filled ghost rows in the band tests and the oracle comparisons of
every kernel at many ragged shapes stand in for memcheck; the staged / ring buffers of the fused kernels are sized by
the same host code that the shape tests drive through every tap remainder.)"""
import ctypes as C

import numpy as np
import pytest

from conftest import rand_image
from image_b200 import FilterType, GaussianBlurParameters
from image_b200 import _lib as L

pytestmark = pytest.mark.gpu
GUARD = 512          # bytes; keeps 16-byte alignment of the destination
DT = {np.dtype(np.uint8): L.U8, np.dtype(np.uint16): L.U16, np.dtype(np.float32): L.F32}


def guarded(cuda, nbytes, call, offset=0):
    """call(dst_ptr) twice: into a guarded buffer (optionally shifted by `offset` bytes) and a plain one"""
    import torch
    buf = torch.full((GUARD + offset + nbytes + GUARD,), 0xA5, dtype=torch.uint8, device=cuda)
    L.check(call(buf.data_ptr() + GUARD + offset))
    plain = torch.zeros(max(nbytes, 1), dtype=torch.uint8, device=cuda)
    L.check(call(plain.data_ptr()))
    torch.cuda.synchronize()
    b = buf.cpu().numpy()
    lo, hi = GUARD + offset, GUARD + offset + nbytes
    assert (b[:lo] == 0xA5).all(), "wrote before the destination"
    assert (b[hi:] == 0xA5).all(), "wrote past the destination"
    assert np.array_equal(b[lo:hi], plain.cpu().numpy()[:nbytes])


@pytest.mark.parametrize("dtype", [np.uint8, np.uint16, np.float32])
@pytest.mark.parametrize("c", [1, 2, 3, 4])
def test_guard_bands(cuda, dtype, c):
    import torch
    lib = L.lib()
    rng = np.random.default_rng(200 + c)
    es = np.dtype(dtype).itemsize
    dt = DT[np.dtype(dtype)]
    st = torch.cuda.current_stream().cuda_stream
    for (h, w) in ((37, 53), (64, 96)):
        img = rand_image(rng, h, w, c, dtype)
        src = torch.from_numpy(img).to(cuda)
        sp = src.data_ptr()
        n = h * w * c * es
        offs = (0, es)  # aligned and element-shifted destinations
        for off in offs:
            for f in (FilterType.Nearest, FilterType.Triangle, FilterType.Lanczos3):
                for (nw, nh) in ((21, 17), (71, 83), (w, 9)):
                    guarded(cuda, nw * nh * c * es, lambda d: lib.b200img_resize_dev(sp, w, h, c, dt, d, nw, nh, int(f), 1.0, 0, st), off)
            for (nw, nh) in ((13, 11), (90, 70), (w // 2, h // 2)):
                guarded(cuda, nw * nh * c * es, lambda d: lib.b200img_thumbnail_dev(sp, w, h, c, dt, d, nw, nh, 0, st), off)
            p = GaussianBlurParameters.new_from_sigma(2.0)
            for path in (L.BLUR_GENERIC, L.BLUR_DYN):
                guarded(cuda, n, lambda d: lib.b200img_gaussian_blur_dev(sp, w, h, c, dt, d, C.byref(p.raw), path, 0, st), off)
            guarded(cuda, n, lambda d: lib.b200img_fast_blur_dev(sp, w, h, c, dt, d, 3.0, st), off)
            k = (C.c_float * 9)(0, -1, 0, -1, 5, -1, 0, -1, 0)
            guarded(cuda, n, lambda d: lib.b200img_filter3x3_dev(sp, w, h, c, dt, d, C.cast(k, C.c_void_p), st), off)
            guarded(cuda, n, lambda d: lib.b200img_unsharpen_dev(sp, w, h, c, dt, d, 1.5, 2, 0, st), off)
            guarded(cuda, n, lambda d: lib.b200img_invert_dev(sp, w, h, c, dt, d, st), off)
            guarded(cuda, n, lambda d: lib.b200img_brighten_dev(sp, w, h, c, dt, d, 20, st), off)
            guarded(cuda, n, lambda d: lib.b200img_contrast_dev(sp, w, h, c, dt, d, 10.0, st), off)
            guarded(cuda, n, lambda d: lib.b200img_huerotate_dev(sp, w, h, c, dt, d, 30, st), off)
            for keep in (0, 1):
                guarded(cuda, h * w * (2 if keep else 1) * es, lambda d: lib.b200img_grayscale_dev(sp, w, h, c, dt, d, keep, st), off)
            for dc, ddt, des in ((3, L.U8, 1), (4, L.F32, 4), (1, L.U16, 2)):
                for mode in (L.CAST, L.CONVERT):
                    guarded(cuda, h * w * dc * des, lambda d: lib.b200img_convert_dev(sp, dt, c, d, ddt, dc, h * w, mode, None, st),
                            off if des <= es else 0)


@pytest.mark.parametrize("dtype", [np.uint8, np.uint16, np.float32])
@pytest.mark.parametrize("c", [1, 3, 4])
def test_guard_bands_fused_kernels_and_bands(cuda, dtype, c):
    """The kernels added in round 2: the fused Gaussian (forced, also where the library would not pick it), the fused
    box blur (u8 rows of whole words, several strips and bands) and the row-band entry points of both blurs."""
    import torch
    lib = L.lib()
    rng = np.random.default_rng(300 + c)
    es = np.dtype(dtype).itemsize
    dt = DT[np.dtype(dtype)]
    st = torch.cuda.current_stream().cuda_stream
    for (h, w) in ((45, 76), (133, 260)):
        img = rand_image(rng, h, w, c, dtype)
        src = torch.from_numpy(img).to(cuda)
        sp = src.data_ptr()
        n = h * w * c * es
        for off in (0, es):
            for sigma in (1.0, 4.0, 10.0):
                p = GaussianBlurParameters.new_from_sigma(sigma)
                for path in (L.BLUR_GENERIC, L.BLUR_DYN):
                    guarded(cuda, n, lambda d: lib.b200img_gaussian_blur_dev(sp, w, h, c, dt, d, C.byref(p.raw), path, L.GAUSS_FUSED, st), off)
                guarded(cuda, n, lambda d: lib.b200img_fast_blur_dev(sp, w, h, c, dt, d, sigma, st), off)
            # a band in the middle of the image from a buffer that holds exactly the rows it reads
            p = GaussianBlurParameters.new_from_sigma(2.0)
            r0, rows = h // 3, h // 3
            s0, sn = C.c_uint32(), C.c_uint32()
            L.check(lib.b200img_gaussian_blur_band_src_rows(h, C.byref(p.raw), r0, rows, C.byref(s0), C.byref(sn)))
            band = torch.from_numpy(np.ascontiguousarray(img[s0.value:s0.value + sn.value])).to(cuda)
            guarded(cuda, rows * w * c * es, lambda d: lib.b200img_gaussian_blur_band_dev(band.data_ptr(), s0.value, sn.value, w, h, c, dt, d, r0, rows,
                                                                                         C.byref(p.raw), L.BLUR_DYN, 0, st), off)
            L.check(lib.b200img_fast_blur_band_src_rows(h, 2.0, r0, rows, C.byref(s0), C.byref(sn)))
            band = torch.from_numpy(np.ascontiguousarray(img[s0.value:s0.value + sn.value])).to(cuda)
            guarded(cuda, rows * w * c * es, lambda d: lib.b200img_fast_blur_band_dev(band.data_ptr(), s0.value, sn.value, w, h, c, dt, d, r0, rows, 2.0, st), off)


def test_guard_bands_fixed_window_resamplers(cuda, monkeypatch):
    """The fixed-window resamplers (resample_int.cu, resample_up2.cu): batches large enough for them to be picked, ragged
    last strips and bands, destinations that are and are not 16-byte aligned."""
    import torch
    monkeypatch.setenv("B200IMG_U2_FORCE", "1")   # the 2x kernel also on jobs smaller than the library would give it
    lib = L.lib()
    rng = np.random.default_rng(400)
    st = torch.cuda.current_stream().cuda_stream
    for c in (1, 3, 4):
        for r, f in ((2, FilterType.Lanczos3), (4, FilterType.CatmullRom), (2, FilterType.Triangle)):
            h, w, n = 92 * r, 372 * r, 12
            imgs = np.stack([rand_image(rng, h, w, c, np.uint8) for _ in range(n)])
            src = torch.from_numpy(imgs).to(cuda)
            nw, nh = w // r, h // r
            # (offsets that keep the kernel choice: the four-channel kernel wants whole-pixel alignment, and the guarded and the
            # plain run are compared bit for bit)
            for off in ((0, 4, 8) if c == 4 else (0, 1, 4)):
                guarded(cuda, n * nw * nh * c, lambda d: lib.b200img_resize_batch_dev(src.data_ptr(), 0, w, h, c, L.U8, d, 0, nw, nh, n, int(f),
                                                                                      1.0, L.FP_FAST, st), off)
    for f in (FilterType.Lanczos3, FilterType.Gaussian):
        h, w, n = 150, 333, 5
        imgs = np.stack([rand_image(rng, h, w, 4, np.float32) for _ in range(n)])
        src = torch.from_numpy(imgs).to(cuda)
        for off in (0, 16):
            guarded(cuda, n * 4 * w * h * 16, lambda d: lib.b200img_resize_batch_dev(src.data_ptr(), 0, w, h, 4, L.F32, d, 0, 2 * w, 2 * h, n, int(f),
                                                                                     1.0, L.FP_FAST, st), off)
