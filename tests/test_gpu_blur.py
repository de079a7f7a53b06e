"""GPU parity: Gaussian blur (f32 engine and Q0.15 engine) and fast_blur against the oracle."""
import numpy as np
import pytest

import oracle as o
from conftest import bench_image, rand_image
from image_b200 import DynamicImage, GaussianBlurParameters, ImageBuffer, imageops
from image_b200 import _lib as L

pytestmark = pytest.mark.gpu
DT = [np.uint8, np.uint16, np.float32]


def _oparams(p):
    r = p.raw
    q = o.GaussParams(r.x_kernel_size, r.x_sigma, r.y_kernel_size, r.y_sigma)
    return q


def _gauss_check(img, p, cuda, host=False):
    import torch
    src = img if host else torch.from_numpy(img).to(cuda)
    get = (lambda t: t) if host else (lambda t: t.cpu().numpy())
    # DynamicImage path: Q0.15 integer for u8/u16 -> bit-exact under either flag
    want_dyn = o.gaussian_blur(img, _oparams(p), "dyn")
    flags_dyn = L.FP_FAST if img.dtype != np.float32 else L.FP_STRICT
    got = get(imageops.blur_dyn(src, p, flags=flags_dyn))
    assert got.tobytes() == want_dyn.tobytes(), f"dyn path differs {img.dtype} {img.shape} {p}"
    for fl in (L.GAUSS_TWO_KERNELS, L.GAUSS_FUSED):   # the row / column kernel pair and the fused kernel, whatever the library would pick
        got = get(imageops.blur_dyn(src, p, flags=flags_dyn | fl))
        assert got.tobytes() == want_dyn.tobytes(), f"dyn path (flag {fl}) differs {img.dtype} {img.shape} {p}"
    # generic path: f32 engine; STRICT bit-exact, FAST within tolerance
    want = o.gaussian_blur(img, _oparams(p), "generic")
    got = get(imageops.blur_advanced(src, p, flags=L.FP_STRICT))
    assert got.tobytes() == want.tobytes(), f"generic strict differs {img.dtype} {img.shape} {p}"
    for fl in (L.FP_FAST, L.FP_FAST | L.GAUSS_TWO_KERNELS, L.FP_FAST | L.GAUSS_FUSED):   # the library's pick, the kernel pair, the fused kernel
        fast = get(imageops.blur_advanced(src, p, flags=fl))
        if img.dtype == np.float32:
            assert np.allclose(fast, want, rtol=1e-5, atol=1e-6)
        else:
            assert np.abs(fast.astype(np.int64) - want.astype(np.int64)).max() <= 1


@pytest.mark.parametrize("dtype", DT)
@pytest.mark.parametrize("c", [1, 2, 3, 4])
def test_gaussian_fused_kernel_shapes(cuda, dtype, c, monkeypatch):
    """The fused kernel's geometry: several strips with a ragged last one, CTA ranges that begin and end inside
    strips (forced CTA counts), every remainder of the 16-tap steps (3 ... 101 taps), images smaller than a strip /
    a chunk, odd row pitches."""
    rng = np.random.default_rng(60 + c)
    img = rand_image(rng, 150, 333 if c < 3 else 131, c, dtype)
    for sigma in (0.5, 0.8, 1.2, 1.5, 2.0, 3.5, 10.0, 15.5):   # 3, 5, 5, 7, 11, 21, 65, 101 taps
        _gauss_check(img, GaussianBlurParameters.new_from_sigma(sigma), cuda)
    for ctas, raw in (("1", "1"), ("7", "0"), ("1000", "1")):   # forced CTA counts; the raw-byte staging path (u8) on and off
        monkeypatch.setenv("B200IMG_GF_CTAS", ctas)
        monkeypatch.setenv("B200IMG_GF_RAW", raw)
        _gauss_check(img, GaussianBlurParameters.new_from_sigma(4.0), cuda)
        _gauss_check(rand_image(rng, 37, 21, c, dtype), GaussianBlurParameters.new_anisotropic_kernel_size(15.0, 41.0), cuda)
    monkeypatch.delenv("B200IMG_GF_CTAS")
    monkeypatch.delenv("B200IMG_GF_RAW")
    _gauss_check(rand_image(rng, 3, 500, c, dtype), GaussianBlurParameters.new_from_sigma(3.0), cuda)
    _gauss_check(rand_image(rng, 300, 2, c, dtype), GaussianBlurParameters.new_from_sigma(3.0), cuda)


def test_gaussian_fused_batch_4k_band_vs_oracle(cuda):
    """The timed configuration itself (8 x 3840x2160 Rgb8, sigma 10, both engines): a 96-row band of the first and
    of the last image against the oracle (which needs the band plus 32 rows either side)."""
    import ctypes as C
    import torch
    lib = L.lib()
    base = bench_image(3840, 2160, 3)
    batch = np.stack([np.roll(base, 37 * i, axis=1) for i in range(8)])
    src = torch.from_numpy(batch).to(cuda)
    dst = torch.empty_like(src)
    p = GaussianBlurParameters.new_from_sigma(10.0)
    stream = torch.cuda.current_stream().cuda_stream
    for path, name in ((L.BLUR_GENERIC, "generic"), (L.BLUR_DYN, "dyn")):
        L.check(lib.b200img_gaussian_blur_batch_dev(src.data_ptr(), 3840, 2160, 3, L.U8, dst.data_ptr(), 8, C.byref(p.raw), path, L.FP_FAST, stream))
        out = dst.cpu().numpy()
        for i, (y0, y1) in ((0, (0, 96)), (7, (2064, 2160)), (3, (1000, 1096))):
            lo, hi = max(0, y0 - 32), min(2160, y1 + 32)
            want = o.gaussian_blur(np.ascontiguousarray(batch[i, lo:hi]), _oparams(p), name)[y0 - lo:y1 - lo]
            d = np.abs(out[i, y0:y1].astype(int) - want.astype(int)).max()
            assert d <= (1 if path == L.BLUR_GENERIC else 0), (name, i, y0, d)


@pytest.mark.parametrize("dtype", DT)
@pytest.mark.parametrize("c", [1, 2, 3, 4])
def test_gaussian_all_types(cuda, dtype, c):
    rng = np.random.default_rng(20 + c)
    img = rand_image(rng, 45, 67, c, dtype)
    for p in (GaussianBlurParameters.new_from_sigma(1.3), GaussianBlurParameters.new_from_radius(4.0),
              GaussianBlurParameters.SMOOTHING_3, GaussianBlurParameters.new_anisotropic_kernel_size(9.0, 3.0)):
        _gauss_check(img, p, cuda)


def test_gaussian_k2_kernels_reduced(cuda):
    """BASELINE config 2 kernels (21 taps = blur(img, 10.0); 65 taps = sigma 10) on a 4K crop."""
    img = bench_image(3840, 2160, 3)[500:900, 1000:1700].copy()
    for p in (GaussianBlurParameters.new_from_radius(10.0), GaussianBlurParameters.new_from_sigma(10.0)):
        _gauss_check(img, p, cuda)


def test_gaussian_kernel_wider_than_image_and_fused_mac(cuda):
    import torch
    rng = np.random.default_rng(31)
    img = rand_image(rng, 5, 9, 3, np.uint8)
    _gauss_check(img, GaussianBlurParameters.new_from_sigma(10.0), cuda)   # 65 taps over a 9x5 image
    _gauss_check(img, GaussianBlurParameters.new_from_sigma(9.0), cuda, host=True)  # >= 55 taps: transient variant
    # multiply_accumulate fused (aarch64 / +fma builds of the reference)
    f = rand_image(rng, 33, 35, 3, np.float32)
    p = GaussianBlurParameters.new_from_sigma(2.0)
    o.set_fused_mac(True)
    try:
        want = o.gaussian_blur(f, _oparams(p), "generic")
    finally:
        o.set_fused_mac(False)
    got = imageops.blur_advanced(torch.from_numpy(f).to(cuda), p, flags=L.FP_STRICT | L.FP_FUSED_MAC).cpu().numpy()
    assert got.tobytes() == want.tobytes()


def test_gaussian_one_tap_quirks(cuda):
    """blur(img, 0.0) (src/imageops/mod.rs:473-478): 1-tap kernels shift right and leave row 0 zero."""
    rng = np.random.default_rng(32)
    for dtype in DT:
        img = rand_image(rng, 7, 8, 2, dtype)
        _gauss_check(img, GaussianBlurParameters.new_from_radius(0.0), cuda)
        _gauss_check(img, GaussianBlurParameters.new_from_sigma(0.05), cuda)
        _gauss_check(img, GaussianBlurParameters.new_anisotropic_kernel_size(1.0, 5.0), cuda)
        _gauss_check(img, GaussianBlurParameters.new_anisotropic_kernel_size(5.0, 1.0), cuda)


def test_gaussian_constant_and_dynimage(cuda):
    img = np.full((40, 50, 4), 201, np.uint8)
    d = DynamicImage(ImageBuffer(img))
    assert (d.blur(3.0).buf.data == 201).all()
    assert (d.blur_advanced(GaussianBlurParameters.new_from_sigma(5.0)).buf.data == 201).all()
    assert (imageops.blur(img, 3.0) == 201).all()


@pytest.mark.parametrize("dtype", DT)
@pytest.mark.parametrize("c", [1, 2, 3, 4])
def test_fast_blur_all_types(cuda, dtype, c):
    import torch
    rng = np.random.default_rng(40 + c)
    img = rand_image(rng, 53, 71, c, dtype)
    for sigma in (0.0, 0.7, 3.0, 10.0):
        want = o.fast_blur(img, sigma)
        got = imageops.fast_blur(torch.from_numpy(img).to(cuda), sigma).cpu().numpy()
        if dtype == np.float32:
            assert np.allclose(got, want, rtol=1e-5, atol=2e-6), sigma
        else:
            assert got.tobytes() == want.tobytes(), (dtype, c, sigma)


@pytest.mark.parametrize("c", [1, 2, 3, 4])
def test_fast_blur_u8_fused_and_unfused_kernels(cuda, c, monkeypatch):
    """u8 images take one fused kernel per box (horizontal rows kept in a shared-memory ring) or, where that does
    not apply, a horizontal and a vertical kernel: both against the oracle, over several strips (ragged last one),
    several bands (forced), windows longer than the image, radii up to the fused kernel's limit and beyond it."""
    import torch
    rng = np.random.default_rng(70 + c)
    shapes = [(150, 600 if c < 4 else 300), (70, 257), (9, 40), (300, 8)]
    for h, w in shapes:
        if (w * c) % 4:
            w += 4 - (w * c) % 4 if c != 3 else 4 - w % 4   # the word-wise kernels need rows of whole words; other widths take the generic ones
        img = rand_image(rng, h, w, c, np.uint8)
        dev = torch.from_numpy(img).to(cuda)
        for sigma in (0.5, 3.0, 10.0, 22.0, 60.0):
            want = o.fast_blur(img, sigma)
            for bands in (None, "1", "5"):
                if bands is None:
                    monkeypatch.delenv("B200IMG_BOX_BANDS", raising=False)
                else:
                    monkeypatch.setenv("B200IMG_BOX_BANDS", bands)
                assert np.array_equal(imageops.fast_blur(dev, sigma).cpu().numpy(), want), (h, w, c, sigma, bands)
            monkeypatch.delenv("B200IMG_BOX_BANDS", raising=False)
            monkeypatch.setenv("B200IMG_BOX_UNFUSED", "1")
            assert np.array_equal(imageops.fast_blur(dev, sigma).cpu().numpy(), want), (h, w, c, sigma, "unfused")
            monkeypatch.delenv("B200IMG_BOX_UNFUSED")


def test_fast_blur_k2_reduced_and_host(cuda):
    img = bench_image(3840, 2160, 3)[300:800, 200:1100].copy()
    assert np.array_equal(imageops.fast_blur(img, 10.0), o.fast_blur(img, 10.0))


def test_fast_blur_large_sigma_and_tiny_images(cuda):
    rng = np.random.default_rng(41)
    for shape in ((1, 1, 3), (1, 17, 1), (23, 1, 4), (6, 5, 2)):
        img = rand_image(rng, *shape, np.uint8)
        for sigma in (1.0, 50.0, 300.0):
            assert np.array_equal(imageops.fast_blur(img, sigma), o.fast_blur(img, sigma)), (shape, sigma)
    assert imageops.fast_blur(np.zeros((0, 0, 3), np.uint8), 3.0).size == 0
    assert np.array_equal(imageops.fast_blur(rand_image(rng, 9, 9, 3, np.uint8), -3.0).shape, (9, 9, 3))


def test_fast_blur_constant_image_property_4k(cuda):
    """fast_blur.rs:535-592 at the K2 size: a constant image stays constant."""
    import torch
    img = torch.full((2160, 3840, 3), 143, dtype=torch.uint8, device=cuda)
    out = imageops.fast_blur(img, 10.0)
    assert int(out.min()) == 143 and int(out.max()) == 143
    out = DynamicImage(ImageBuffer(img)).blur_advanced(GaussianBlurParameters.new_from_sigma(10.0)).buf.data
    assert int(out.min()) == 143 and int(out.max()) == 143


def test_host_batch_entry_points_match_single_calls(cuda):
    """b200img_*_batch (host-resident, pipelined on three streams) == the per-image host calls."""
    import ctypes as C
    lib = L.lib()
    rng = np.random.default_rng(50)
    n, h, w, c = 7, 61, 90, 3  # more images than pipeline slots
    batch = rng.integers(0, 256, (n, h, w, c), dtype=np.uint8)
    p = GaussianBlurParameters.new_from_sigma(2.2)
    out = np.empty_like(batch)
    L.check(lib.b200img_gaussian_blur_batch(batch.ctypes.data, w, h, c, L.U8, out.ctypes.data, n, C.byref(p.raw), L.BLUR_GENERIC, L.FP_FAST))
    for i in range(n):
        assert np.array_equal(out[i], imageops.blur_advanced(batch[i], p)), i
    L.check(lib.b200img_fast_blur_batch(batch.ctypes.data, w, h, c, L.U8, out.ctypes.data, n, 3.0))
    for i in range(n):
        assert np.array_equal(out[i], o.fast_blur(batch[i], 3.0)), i
    small = np.empty((n, 20, 33, c), np.uint8)
    L.check(lib.b200img_resize_batch(batch.ctypes.data, w, h, c, L.U8, small.ctypes.data, 33, 20, n, 4, 1.0, L.FP_STRICT))
    for i in range(n):
        assert np.array_equal(small[i], o.resize(batch[i], 33, 20, "lanczos3")), i
    # a second, larger call re-sizes the staging buffers
    big = rng.integers(0, 256, (4, 200, 300, 4), dtype=np.uint8)
    outb = np.empty_like(big)
    L.check(lib.b200img_fast_blur_batch(big.ctypes.data, 300, 200, 4, L.U8, outb.ctypes.data, 4, 5.0))
    for i in range(4):
        assert np.array_equal(outb[i], o.fast_blur(big[i], 5.0)), i


@pytest.mark.parametrize("dtype", DT)
def test_long_kernels(cuda, dtype):
    """benches/blur.rs uses sigma = 50 (331 taps): the tiled kernels take kernels up to 384 taps, the
    direct kernels everything beyond (sigma = 60: 397 taps) and tiles that do not fit shared memory"""
    rng = np.random.default_rng(31)
    for c in (1, 3, 4):
        img = rand_image(rng, 70, 90, c, dtype)
        _gauss_check(img, GaussianBlurParameters.new_from_sigma(50.0), cuda)
    img = rand_image(rng, 40, 50, 3, dtype)
    _gauss_check(img, GaussianBlurParameters.new_from_sigma(60.0), cuda)
    _gauss_check(img, GaussianBlurParameters.new_anisotropic_kernel_size(301.0, 5.0), cuda)
