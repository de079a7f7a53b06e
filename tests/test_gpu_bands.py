"""One image over several devices through the library's own band calls (csrc/multi.cu): the stitched
result must be the whole-image result, seams included.  A device ordinal may repeat, so the whole
mechanism (plan, ghost-row copies, per-band kernels) also runs on a 1-GPU box; with two or more GPUs
the bands alternate between devices 0 and 1 and the ghost rows cross NVLink."""
import numpy as np
import pytest

import oracle as o
from image_b200 import FilterType, GaussianBlurParameters, imageops
from image_b200 import _lib as L
from image_b200.multi import run_bands

pytestmark = pytest.mark.gpu


def _device_sets():
    import torch
    sets = [[0], [0, 0, 0], [0] * 7]
    if torch.cuda.device_count() >= 2:
        sets += [[0, 1], [0, 1, 0, 1, 1]]
    return sets


@pytest.mark.parametrize("filt,name", [(FilterType.Lanczos3, "lanczos3"), (FilterType.CatmullRom, "catmullrom"),
                                        (FilterType.Nearest, "nearest"), (FilterType.Triangle, "triangle")])
def test_resize_bands_bit_exact(cuda, filt, name):
    rng = np.random.default_rng(3)
    for dtype, (h, w, c), (nh, nw) in ((np.float32, (97, 61, 4), (194, 122)), (np.uint8, (203, 88, 3), (67, 41)),
                                       (np.uint16, (64, 50, 2), (64, 75)), (np.uint8, (40, 33, 4), (40, 33))):
        img = rng.random((h, w, c), dtype=np.float32) if dtype == np.float32 else rng.integers(0, np.iinfo(dtype).max + 1, (h, w, c), dtype=dtype)
        want = o.resize(img, nw, nh, name)
        for devs in _device_sets():
            got = run_bands("resize", img, devs, nw=nw, nh=nh, filter=filt, flags=L.FP_STRICT)
            assert got.tobytes() == want.tobytes(), (name, dtype, devs)


def test_gaussian_bands(cuda):
    rng = np.random.default_rng(4)
    for sigma in (0.1, 1.5, 6.0):   # 0.1: the one-tap kernel (no halo; its row-0 quirk belongs to the image's row 0 only)
        p = GaussianBlurParameters.new_from_sigma(sigma)
        op = o.gauss_params_from_sigma(sigma)
        for dtype in (np.uint8, np.uint16):
            img = rng.integers(0, np.iinfo(dtype).max + 1, (131, 77, 3), dtype=dtype)
            want = o.gaussian_blur(img, op, "dyn")   # Q0.15: bit-exact, seams included
            for devs in _device_sets():
                got = run_bands("gaussian", img, devs, params=p, path=L.BLUR_DYN)
                assert got.tobytes() == want.tobytes(), (sigma, dtype, devs)
        img = rng.random((90, 45, 4), dtype=np.float32)
        want = o.gaussian_blur(img, op, "generic")
        for devs in _device_sets():
            got = run_bands("gaussian", img, devs, params=p, path=L.BLUR_GENERIC, flags=L.FP_STRICT)
            assert got.tobytes() == want.tobytes(), (sigma, devs)
            fast = run_bands("gaussian", img, devs, params=p, path=L.BLUR_GENERIC, flags=L.FP_FAST)
            assert np.allclose(fast, want, rtol=1e-5, atol=1e-6)


def test_gaussian_bands_large_through_the_fused_row_window(cuda):
    """Bands of a large image with a long kernel run the fused kernel on exactly the band's rows (no scratch, no
    copy): the stitched result equals the whole-image result bit for bit (Q0.15) / the oracle within +-1 (f32 engine)."""
    rng = np.random.default_rng(6)
    img = rng.integers(0, 256, (2100, 1536, 3), dtype=np.uint8)
    p = GaussianBlurParameters.new_from_sigma(10.0)
    whole_q = imageops.blur_dyn(img, p)
    whole_f = imageops.blur_advanced(img, p)
    for devs in _device_sets()[1:]:
        assert np.array_equal(run_bands("gaussian", img, devs, params=p, path=L.BLUR_DYN), whole_q), devs
        assert np.array_equal(run_bands("gaussian", img, devs, params=p, path=L.BLUR_GENERIC), whole_f), devs
    band = slice(690, 760)   # across the first seam of the three-band split
    want = o.gaussian_blur(np.ascontiguousarray(img[band.start - 32:band.stop + 32]), o.gauss_params_from_sigma(10.0), "dyn")[32:-32]
    assert np.array_equal(whole_q[band], want)


def test_fast_blur_bands(cuda):
    rng = np.random.default_rng(5)
    for sigma in (1.0, 4.0):   # sigma 4: radii 3+3+5 = 11 ghost rows per side; 7 bands of 23 rows reach two bands away
        for dtype in (np.uint8, np.uint16):
            img = rng.integers(0, np.iinfo(dtype).max + 1, (161, 59, 3), dtype=dtype)
            want = o.fast_blur(img, sigma)
            for devs in _device_sets():
                got = run_bands("fast_blur", img, devs, sigma=sigma)
                assert got.tobytes() == want.tobytes(), (sigma, dtype, devs)
        # f32 running sums depend on where a band starts: same tolerance as against the oracle elsewhere
        img = rng.random((120, 40, 4), dtype=np.float32)
        want = o.fast_blur(img, sigma)
        got = run_bands("fast_blur", img, [0, 0, 0], sigma=sigma)
        assert np.allclose(got, want, rtol=1e-4, atol=1e-5)


def test_band_errors(cuda):
    import torch
    lib = L.lib()
    bands = (L.Band * 2)()
    L.check(lib.b200img_plan_fast_blur_bands(100, 2.0, 2, bands))
    buf = torch.zeros(1 << 20, dtype=torch.uint8, device=cuda)
    for b in bands:
        b.device, b.d_src, b.d_dst = 0, buf.data_ptr(), buf.data_ptr()
    bands[1].device = 99
    assert lib.b200img_fast_blur_bands(bands, 2, 16, 100, 3, L.U8, 2.0, None) == L.ERR_BAD_ARG
    bands[1].device = 0
    bands[0].own_rows = 10   # rows 10..49 now belong to nobody, and band 1 reads some of them
    assert lib.b200img_fast_blur_bands(bands, 2, 16, 100, 3, L.U8, 2.0, None) == L.ERR_BAD_ARG
    assert "no band owns" in L.last_error()
