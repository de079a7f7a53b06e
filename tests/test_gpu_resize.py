"""GPU parity: the resampler through the C ABI against the oracle.
STRICT flag: bit-exact.  FAST flag (FMA): +-1 LSB for u8/u16, 1e-5 relative for f32."""
import numpy as np
import pytest

import oracle as o
from conftest import bench_image, rand_image
from image_b200 import FilterType, imageops
from image_b200 import _lib as L

pytestmark = pytest.mark.gpu
FILTERS = list(FilterType)


def _check(img, nw, nh, f, cuda, acc_init=1.0, host=False, unfused=False, tiled=False):
    import torch
    want = o.resize(img, nw, nh, int(f), acc_init=acc_init)
    extra = (L.RESIZE_UNFUSED if unfused else 0) | (L.RESIZE_TILED if tiled else 0)
    src = img if host else torch.from_numpy(img).to(cuda)
    strict = imageops.resize(src, nw, nh, f, acc_init=acc_init, flags=L.FP_STRICT | extra)
    fast = imageops.resize(src, nw, nh, f, acc_init=acc_init, flags=L.FP_FAST | extra)
    if not host:
        strict, fast = strict.cpu().numpy(), fast.cpu().numpy()
    assert strict.shape == want.shape
    assert strict.tobytes() == want.tobytes(), f"strict differs: {f.name} {img.shape}->{(nh, nw)} max|d|={np.abs(strict.astype(np.float64) - want).max()}"
    if img.dtype == np.float32:
        assert np.allclose(fast, want, rtol=1e-5, atol=1e-6)
    else:
        assert np.abs(fast.astype(np.int64) - want.astype(np.int64)).max() <= 1
    return fast, want


@pytest.mark.parametrize("f", FILTERS)
@pytest.mark.parametrize("dtype", [np.uint8, np.uint16, np.float32])
@pytest.mark.parametrize("c", [1, 2, 3, 4])
def test_resize_down_up_all_types(cuda, f, dtype, c):
    rng = np.random.default_rng(c * 10 + int(f))
    img = rand_image(rng, 61, 83, c, dtype)
    for acc in (1.0, 0.0):
        _check(img, 40, 29, f, cuda, acc_init=acc)     # down
        _check(img, 131, 97, f, cuda, acc_init=acc)    # up
    _check(img, 83, 20, f, cuda)                       # height only (horizontal pass still runs)
    _check(img, 30, 61, f, cuda)                       # width only


@pytest.mark.parametrize("f", FILTERS)
def test_k1_shape(cuda, f):
    """BASELINE config 1: 1920x1080 Rgba8 -> 960x540 (bench generator + alpha)."""
    img = bench_image(1920, 1080, 4)
    fast, want = _check(img, 960, 540, f, cuda)
    if f == FilterType.Nearest:
        assert np.array_equal(fast, want)


@pytest.mark.parametrize("f", [FilterType.Nearest, FilterType.Triangle, FilterType.Lanczos3])
def test_bench_shapes_rgb8(cuda, f):
    """benches/imageops.rs:95-124: 1920x1080 Rgb8 -> 400x300 and (reduced) upscale."""
    img = bench_image(1920, 1080, 3)
    _check(img, 400, 300, f, cuda)
    _check(img[:300, :400].copy(), 1000, 750, f, cuda)


def test_k4_batch(cuda):
    """BASELINE config 4 at a reduced batch: 1024x1024 Rgba8 -> 256x256 CatmullRom, one launch."""
    import torch
    rng = np.random.default_rng(4)
    batch = rng.integers(0, 256, (3, 1024, 1024, 4), dtype=np.uint8)
    got = imageops.resize_batch(torch.from_numpy(batch).to(cuda), 256, 256, FilterType.CatmullRom, flags=L.FP_STRICT).cpu().numpy()
    for i in range(3):
        assert np.array_equal(got[i], o.resize(batch[i], 256, 256, "catmullrom")), i


def test_k5_band_f32_upscale(cuda):
    """BASELINE config 5 on a crop: Rgba32F Lanczos3 2x upscale, literal and corrected modes."""
    rng = np.random.default_rng(5)
    img = rng.random((96, 160, 4), dtype=np.float32)
    _check(img, 320, 192, FilterType.Lanczos3, cuda, acc_init=1.0)
    fast, want = _check(img, 320, 192, FilterType.Lanczos3, cuda, acc_init=0.0)
    assert want.min() >= 0 and want.max() <= 1 and want.std() > 0.1


def test_band_entry_point(cuda):
    """Row-band split: two bands with their halos reproduce the whole image bit for bit."""
    import ctypes as C
    import torch
    rng = np.random.default_rng(6)
    img = rng.random((120, 70, 4), dtype=np.float32)
    nw, nh = 140, 240
    want = o.resize(img, nw, nh, "lanczos3", acc_init=0.0)
    lib = L.lib()
    dev = torch.from_numpy(img).to(cuda)
    out = torch.empty((nh, nw, 4), dtype=torch.float32, device=cuda)
    for r0, rows in ((0, 100), (100, 140)):
        s0, sn = C.c_uint32(), C.c_uint32()
        L.check(lib.b200img_resize_band_src_rows(120, nh, 4, r0, rows, C.byref(s0), C.byref(sn)))
        band = dev[s0.value:s0.value + sn.value].contiguous()
        dst = torch.empty((rows, nw, 4), dtype=torch.float32, device=cuda)
        L.check(lib.b200img_resize_band_dev(band.data_ptr(), s0.value, sn.value, 70, 120, 4, L.F32, dst.data_ptr(),
                                            r0, rows, nw, nh, 4, 0.0, L.FP_STRICT, None))
        out[r0:r0 + rows] = dst
    assert out.cpu().numpy().tobytes() == want.tobytes()
    # a band that lacks its halo is rejected
    s0, sn = C.c_uint32(), C.c_uint32()
    L.check(lib.b200img_resize_band_src_rows(120, nh, 4, 100, 140, C.byref(s0), C.byref(sn)))
    assert lib.b200img_resize_band_dev(dev.data_ptr(), s0.value + 1, sn.value - 1, 70, 120, 4, L.F32, out.data_ptr(),
                                       100, 140, nw, nh, 4, 0.0, 0, None) == L.ERR_BAD_ARG


@pytest.mark.parametrize("f", FILTERS)
@pytest.mark.parametrize("dtype", [np.uint8, np.uint16])
def test_tile_kernel_on_downscales(cuda, f, dtype):
    """Downscales of integer images default to the streaming kernel; the tile kernel must agree."""
    rng = np.random.default_rng(21)
    for c in (1, 3, 4):
        img = rand_image(rng, 90, 140, c, dtype)
        _check(img, 61, 37, f, cuda, tiled=True)
        _check(img, 33, 20, f, cuda, tiled=True, acc_init=0.0)


def test_streaming_kernel_shapes(cuda):
    """Strips, bands and ragged edges of the streaming kernel: many bands, several strips, odd sizes."""
    rng = np.random.default_rng(22)
    for (h, w, c, nh, nw, f) in [(700, 900, 4, 233, 301, FilterType.Lanczos3), (513, 1027, 3, 100, 257, FilterType.CatmullRom),
                                 (1024, 1024, 4, 256, 256, FilterType.CatmullRom), (300, 2000, 1, 299, 500, FilterType.Triangle),
                                 (257, 64, 2, 17, 63, FilterType.Gaussian), (1080, 1920, 4, 540, 960, FilterType.Lanczos3)]:
        _check(rand_image(rng, h, w, c, np.uint8), nw, nh, f, cuda)
    _check(rand_image(rng, 400, 333, 4, np.uint16), 111, 160, FilterType.Lanczos3, cuda)
    _check(rand_image(rng, 400, 334, 3, np.uint16), 112, 90, FilterType.Triangle, cuda, acc_init=0.0)


@pytest.mark.parametrize("f", FILTERS)
def test_unfused_structure_matches(cuda, f):
    rng = np.random.default_rng(8)
    img = rand_image(rng, 50, 64, 3, np.uint8)
    _check(img, 31, 23, f, cuda, unfused=True)
    _check(rand_image(rng, 33, 20, 4, np.float32), 77, 50, f, cuda, unfused=True)


def test_extreme_downscale_falls_back(cuda):
    """bug_1600 (src/imageops/sample.rs:1892-1897): 629x627 -> 22x22 Lanczos3 is not all zero."""
    img = np.full((627, 629, 4), 255, np.uint8)
    fast, want = _check(img, 22, 22, FilterType.Lanczos3, cuda)
    assert fast.any()
    rng = np.random.default_rng(12)
    _check(rand_image(rng, 900, 700, 3, np.uint8), 5, 3, FilterType.Gaussian, cuda)


def test_edge_cases(cuda):
    import torch
    rng = np.random.default_rng(9)
    img = rand_image(rng, 9, 11, 3, np.uint8)
    # same size: copy (sample.rs:987-991)
    assert np.array_equal(imageops.resize(torch.from_numpy(img).to(cuda), 11, 9, FilterType.Lanczos3).cpu().numpy(), img)
    # empty source: zero-filled destination (977-984; issue_2340)
    for shape in ((0, 5, 1), (5, 0, 1)):
        z = imageops.resize(torch.zeros(shape, dtype=torch.uint8, device=cuda), 16, 16, FilterType.Lanczos3)
        assert z.shape == (16, 16, 1) and not z.any()
    # empty destination
    assert imageops.resize(img, 0, 7, FilterType.Triangle).shape == (7, 0, 3)
    # 1x1 source, 1-pixel destination, host-pointer path
    _check(rand_image(rng, 1, 1, 4, np.uint16), 5, 3, FilterType.CatmullRom, cuda, host=True)
    _check(img, 1, 1, FilterType.Lanczos3, cuda, host=True)
    _check(rand_image(rng, 3, 200, 1, np.uint8), 333, 2, FilterType.Triangle, cuda, host=True)


def test_resize_transparent_image_invariant(cuda):
    """sample.rs:1861-1897: alpha never 253/254 in the centre, all filters (literal mode)."""
    img = np.zeros((32, 32, 4), np.uint8)
    img[4:28, 4:28] = [200, 100, 50, 255]
    for f in FILTERS:
        a = imageops.resize(img, 16, 16, f)[5:11, 5:11, 3]
        assert not np.isin(a, [253, 254]).any(), f


def test_nearest_roundtrip_identity(cuda):
    """tests/regression.rs:154-174: Nearest 2x up then down is the identity (acc_init = 0)."""
    rng = np.random.default_rng(2)
    img = rand_image(rng, 66, 129, 4, np.uint8)
    up = imageops.resize(img, 258, 132, FilterType.Nearest, acc_init=0.0)
    assert np.array_equal(imageops.resize(up, 129, 66, FilterType.Nearest, acc_init=0.0), img)


def test_large_image_linearity_property(cuda):
    """Size-independent property on a large image (no oracle): resize of a constant image is the
    constant (+acc_init), and Nearest output pixels are source pixels."""
    import torch
    img = torch.full((2160, 3840, 4), 77, dtype=torch.uint8, device=cuda)
    out = imageops.resize(img, 1920, 1080, FilterType.Lanczos3, acc_init=0.0)
    assert int(out.min()) >= 76 and int(out.max()) <= 78
    g = torch.Generator(device=cuda).manual_seed(3)
    rnd = torch.randint(0, 256, (2160, 3840, 4), dtype=torch.uint8, device=cuda, generator=g)
    near = imageops.resize(rnd, 960, 540, FilterType.Nearest, acc_init=0.0)
    l_v, _, _ = imageops.build_weights(FilterType.Nearest, 2160, 540, 0)
    l_h, _, _ = imageops.build_weights(FilterType.Nearest, 3840, 960, 1)
    idx_v = torch.from_numpy(l_v.astype(np.int64)).to(cuda)
    idx_h = torch.from_numpy(l_h.astype(np.int64)).to(cuda)
    assert torch.equal(near, rnd[idx_v][:, idx_h])


def _int_ratio_check(img, r, f, cuda, acc_init=1.0):
    """u8 downscale by the even integer ratio r: the fixed-window kernel (default) and the table-driven kernels
    (RESIZE_GENERAL) are both within +-1 LSB of the oracle."""
    import torch
    h, w = img.shape[:2]
    nw, nh = w // r, h // r
    want = o.resize(img, nw, nh, int(f), acc_init=acc_init).astype(np.int64)
    src = torch.from_numpy(img).to(cuda)
    fast = imageops.resize(src, nw, nh, f, acc_init=acc_init, flags=L.FP_FAST).cpu().numpy().astype(np.int64)
    general = imageops.resize(src, nw, nh, f, acc_init=acc_init, flags=L.FP_FAST | L.RESIZE_GENERAL).cpu().numpy().astype(np.int64)
    d = np.abs(fast - want)
    assert d.max() <= 1, f"{f.name} r={r} {img.shape}: max|d|={d.max()} at {np.argwhere(d > 1)[:5].tolist()}"
    assert np.abs(general - want).max() <= 1
    return float((d > 0).mean())


@pytest.mark.parametrize("f", [FilterType.Triangle, FilterType.CatmullRom, FilterType.Gaussian, FilterType.Lanczos3])
@pytest.mark.parametrize("c", [1, 2, 3, 4])
@pytest.mark.parametrize("r", [2, 4])
def test_resize_integer_ratio_u8(cuda, f, c, r):
    """Even integer ratios run the fixed-window kernel (resample_int.cu): every filter, channel count, ragged strip and
    band ends, images smaller than a window, and both accumulator start values (sample.rs:547)."""
    rng = np.random.default_rng(100 * r + 10 * c + int(f))
    off = 0.0
    for (w, h) in [(1920, 1080), (r * 200 + 8 * r, r * 67), (r * 36, r * 500), (8 * r, 4 * r)]:
        if (w * c) % 4:
            w += 4 * r
        img = rand_image(rng, h, w, c, np.uint8)
        off = max(off, _int_ratio_check(img, r, f, cuda))
    # extremes: a saturated image stays saturated (weights sum to one up to rounding), black stays black
    for v in (0, 255):
        img = np.full((r * 64, r * 320, c), v, np.uint8)
        _int_ratio_check(img, r, f, cuda, acc_init=0.0)
    assert off < 0.02  # +-1 LSB differences are rounding ties, not a bias


def test_resize_integer_ratio_batch_and_bands(cuda):
    """The batch entry point and the row-band entry point on the fixed-window kernel."""
    import torch
    rng = np.random.default_rng(77)
    imgs = np.stack([rand_image(rng, 384, 512, 4, np.uint8) for _ in range(24)])
    src = torch.from_numpy(imgs).to(cuda)
    dst = torch.empty((24, 96, 128, 4), dtype=torch.uint8, device=cuda)
    st = torch.cuda.current_stream().cuda_stream
    lib = L.lib()
    L.check(lib.b200img_resize_batch_dev(src.data_ptr(), 0, 512, 384, 4, L.U8, dst.data_ptr(), 0, 128, 96, 24,
                                         int(FilterType.CatmullRom), 1.0, L.FP_FAST, st))
    got = dst.cpu().numpy()
    for i in (0, 11, 23):
        want = o.resize(imgs[i], 128, 96, int(FilterType.CatmullRom))
        assert np.abs(got[i].astype(np.int64) - want.astype(np.int64)).max() <= 1
    # bands of one large image: each band reads only the rows its outputs tap
    img = rand_image(rng, 4000, 1024, 3, np.uint8)
    want = o.resize(img, 512, 2000, int(FilterType.Lanczos3)).astype(np.int64)
    out = np.zeros((2000, 512, 3), np.uint8)
    import ctypes
    for r0, rows in [(0, 700), (700, 601), (1301, 699)]:
        s0, sn = ctypes.c_uint32(), ctypes.c_uint32()
        L.check(lib.b200img_resize_band_src_rows(4000, 2000, int(FilterType.Lanczos3), r0, rows, ctypes.byref(s0), ctypes.byref(sn)))
        band = torch.from_numpy(img[s0.value:s0.value + sn.value].copy()).to(cuda)
        d = torch.empty((rows, 512, 3), dtype=torch.uint8, device=cuda)
        L.check(lib.b200img_resize_band_dev(band.data_ptr(), s0.value, sn.value, 1024, 4000, 3, L.U8, d.data_ptr(), r0, rows,
                                            512, 2000, int(FilterType.Lanczos3), 1.0, L.FP_FAST, st))
        out[r0:r0 + rows] = d.cpu().numpy()
    assert np.abs(out.astype(np.int64) - want).max() <= 1


@pytest.mark.parametrize("f", [FilterType.Triangle, FilterType.CatmullRom, FilterType.Lanczos3, FilterType.Gaussian])
def test_resize_up2_rgba32f(cuda, f, monkeypatch):
    """Exact 2x upscales of Rgba32F with a support-3 filter run the register-window kernel (resample_up2.cu): ragged strips
    and bands, both accumulator start values, a batch, and row bands that start and end on odd destination rows."""
    import ctypes
    import torch
    monkeypatch.setenv("B200IMG_U2_FORCE", "1")   # the library keeps this kernel for jobs of 6 M source pixels and more
    rng = np.random.default_rng(int(f))
    lib = L.lib()
    st = torch.cuda.current_stream().cuda_stream
    for (h, w) in [(300, 517), (1031, 250), (64, 2048)]:
        img = rand_image(rng, h, w, 4, np.float32)
        for acc in (1.0, 0.0):
            want = o.resize(img, 2 * w, 2 * h, int(f), acc_init=acc)
            src = torch.from_numpy(img).to(cuda)
            got = imageops.resize(src, 2 * w, 2 * h, f, acc_init=acc, flags=L.FP_FAST).cpu().numpy()
            general = imageops.resize(src, 2 * w, 2 * h, f, acc_init=acc, flags=L.FP_FAST | L.RESIZE_GENERAL).cpu().numpy()
            assert np.allclose(got, want, rtol=1e-5, atol=1e-6), f"{f.name} {img.shape}: {np.abs(got - want).max()}"
            assert np.allclose(general, want, rtol=1e-5, atol=1e-6)
    # batch
    imgs = np.stack([rand_image(rng, 200, 320, 4, np.float32) for _ in range(6)])
    src = torch.from_numpy(imgs).to(cuda)
    dst = torch.empty((6, 400, 640, 4), dtype=torch.float32, device=cuda)
    L.check(lib.b200img_resize_batch_dev(src.data_ptr(), 0, 320, 200, 4, L.F32, dst.data_ptr(), 0, 640, 400, 6, int(f), 1.0, L.FP_FAST, st))
    got = dst.cpu().numpy()
    for i in (0, 5):
        assert np.allclose(got[i], o.resize(imgs[i], 640, 400, int(f)), rtol=1e-5, atol=1e-6)
    # bands with odd seams
    img = rand_image(rng, 900, 400, 4, np.float32)
    want = o.resize(img, 800, 1800, int(f))
    out = np.zeros_like(want)
    for r0, rows in [(0, 601), (601, 598), (1199, 601)]:
        s0, sn = ctypes.c_uint32(), ctypes.c_uint32()
        L.check(lib.b200img_resize_band_src_rows(900, 1800, int(f), r0, rows, ctypes.byref(s0), ctypes.byref(sn)))
        band = torch.from_numpy(img[s0.value:s0.value + sn.value].copy()).to(cuda)
        d = torch.empty((rows, 800, 4), dtype=torch.float32, device=cuda)
        L.check(lib.b200img_resize_band_dev(band.data_ptr(), s0.value, sn.value, 400, 900, 4, L.F32, d.data_ptr(), r0, rows,
                                            800, 1800, int(f), 1.0, L.FP_FAST, st))
        out[r0:r0 + rows] = d.cpu().numpy()
    assert np.allclose(out, want, rtol=1e-5, atol=1e-6)


def test_resize_fixed_window_fuzz(cuda, monkeypatch):
    """Random shapes through the fixed-window kernels (batches, so that they are picked): u8 downscales by 2 and 4 and
    Rgba32F 2x upscales against the oracle on the first and last image."""
    import torch
    monkeypatch.setenv("B200IMG_U2_FORCE", "1")
    rng = np.random.default_rng(2024)
    lib = L.lib()
    st = torch.cuda.current_stream().cuda_stream
    filters = [FilterType.Triangle, FilterType.CatmullRom, FilterType.Gaussian, FilterType.Lanczos3]
    for it in range(10):
        r = int(rng.choice([2, 4]))
        c = int(rng.integers(1, 5))
        nw, nh = int(rng.integers(40, 400)), int(rng.integers(30, 300))
        while (nw * r * c) % 4:
            nw += 1
        f = filters[int(rng.integers(0, 4))]
        n = int(rng.integers(6, 20))
        acc = float(rng.choice([0.0, 1.0]))
        imgs = rng.integers(0, 256, (n, nh * r, nw * r, c), dtype=np.uint8)
        src = torch.from_numpy(imgs).to(cuda)
        dst = torch.empty((n, nh, nw, c), dtype=torch.uint8, device=cuda)
        L.check(lib.b200img_resize_batch_dev(src.data_ptr(), 0, nw * r, nh * r, c, L.U8, dst.data_ptr(), 0, nw, nh, n, int(f), acc, L.FP_FAST, st))
        got = dst.cpu().numpy().astype(np.int64)
        for i in (0, n - 1):
            want = o.resize(imgs[i], nw, nh, int(f), acc_init=acc).astype(np.int64)
            assert np.abs(got[i] - want).max() <= 1, (it, r, c, nw, nh, f.name, n)
    for it in range(4):
        w, h = int(rng.integers(60, 500)), int(rng.integers(40, 400))
        f = [FilterType.Lanczos3, FilterType.Gaussian, FilterType.CatmullRom, FilterType.Triangle][it & 3]
        n = int(rng.integers(3, 8))
        imgs = rng.random((n, h, w, 4), dtype=np.float32)
        src = torch.from_numpy(imgs).to(cuda)
        dst = torch.empty((n, 2 * h, 2 * w, 4), dtype=torch.float32, device=cuda)
        L.check(lib.b200img_resize_batch_dev(src.data_ptr(), 0, w, h, 4, L.F32, dst.data_ptr(), 0, 2 * w, 2 * h, n, int(f), 1.0, L.FP_FAST, st))
        got = dst.cpu().numpy()
        for i in (0, n - 1):
            want = o.resize(imgs[i], 2 * w, 2 * h, int(f))
            assert np.allclose(got[i], want, rtol=1e-5, atol=1e-6), (it, w, h, f.name)
    # non-finite samples spread exactly as far as in the reference (the kernel multiplies the same seven taps per axis, the
    # zero-weight end included, and pads with finite zeros): a batch large enough for the 2x kernel to be picked
    imgs = rng.random((8, 160, 250, 4), dtype=np.float32) * 0.5 + 0.1
    imgs[0, 80, 125] = np.inf
    imgs[0, 3, 5, 1] = np.nan
    imgs[7, 159, 249, 2] = np.nan
    src = torch.from_numpy(imgs).to(cuda)
    dst = torch.empty((8, 320, 500, 4), dtype=torch.float32, device=cuda)
    L.check(lib.b200img_resize_batch_dev(src.data_ptr(), 0, 250, 160, 4, L.F32, dst.data_ptr(), 0, 500, 320, 8, int(FilterType.Lanczos3), 0.0,
                                         L.FP_FAST, st))
    got = dst.cpu().numpy()
    for i in (0, 7):
        want = o.resize(imgs[i], 500, 320, int(FilterType.Lanczos3), acc_init=0.0)
        assert np.array_equal(got[i] == 0, want == 0)          # NaN -> 0 in the f32 clamp (primitive_sealed.rs:127-137)
        assert np.allclose(got[i], want, rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("dtype", [np.uint8, np.uint16])
@pytest.mark.parametrize("f", [FilterType.Triangle, FilterType.CatmullRom, FilterType.Lanczos3, FilterType.Gaussian])
def test_resize_up2_rgba_integer_types(cuda, f, dtype, monkeypatch):
    """The 2x upscale kernel on Rgba8 / Rgba16 (same passes, integer load and store): +-1 LSB of the oracle, ragged strips and
    bands, a batch, odd-seamed row bands, saturated and black images."""
    import ctypes
    import torch
    monkeypatch.setenv("B200IMG_U2_FORCE", "1")
    rng = np.random.default_rng(31 + int(f))
    lib = L.lib()
    st = torch.cuda.current_stream().cuda_stream
    dt = L.U8 if dtype == np.uint8 else L.U16
    top = np.iinfo(dtype).max
    for (h, w) in [(300, 516), (1031, 252), (64, 2048)]:
        img = rand_image(rng, h, w, 4, dtype)
        for acc in (1.0, 0.0):
            want = o.resize(img, 2 * w, 2 * h, int(f), acc_init=acc).astype(np.int64)
            got = imageops.resize(torch.from_numpy(img).to(cuda), 2 * w, 2 * h, f, acc_init=acc, flags=L.FP_FAST).cpu().numpy().astype(np.int64)
            assert np.abs(got - want).max() <= 1, (f.name, img.shape, acc)
    for v in (0, top):
        img = np.full((200, 320, 4), v, dtype)
        got = imageops.resize(torch.from_numpy(img).to(cuda), 640, 400, f, acc_init=0.0, flags=L.FP_FAST).cpu().numpy().astype(np.int64)
        assert np.abs(got - o.resize(img, 640, 400, int(f), acc_init=0.0).astype(np.int64)).max() <= 1
    imgs = np.stack([rand_image(rng, 200, 320, 4, dtype) for _ in range(6)])
    src = torch.from_numpy(imgs).to(cuda)
    dst = torch.empty((6, 400, 640, 4), dtype=src.dtype, device=cuda)
    L.check(lib.b200img_resize_batch_dev(src.data_ptr(), 0, 320, 200, 4, dt, dst.data_ptr(), 0, 640, 400, 6, int(f), 1.0, L.FP_FAST, st))
    got = dst.cpu().numpy().astype(np.int64)
    for i in (0, 5):
        assert np.abs(got[i] - o.resize(imgs[i], 640, 400, int(f)).astype(np.int64)).max() <= 1
    img = rand_image(rng, 900, 400, 4, dtype)
    want = o.resize(img, 800, 1800, int(f)).astype(np.int64)
    out = np.zeros((1800, 800, 4), dtype)
    for r0, rows in [(0, 601), (601, 598), (1199, 601)]:
        s0, sn = ctypes.c_uint32(), ctypes.c_uint32()
        L.check(lib.b200img_resize_band_src_rows(900, 1800, int(f), r0, rows, ctypes.byref(s0), ctypes.byref(sn)))
        band = torch.from_numpy(img[s0.value:s0.value + sn.value].copy()).to(cuda)
        d = torch.empty((rows, 800, 4), dtype=band.dtype, device=cuda)
        L.check(lib.b200img_resize_band_dev(band.data_ptr(), s0.value, sn.value, 400, 900, 4, dt, d.data_ptr(), r0, rows,
                                            800, 1800, int(f), 1.0, L.FP_FAST, st))
        out[r0:r0 + rows] = d.cpu().numpy()
    assert np.abs(out.astype(np.int64) - want).max() <= 1


@pytest.mark.parametrize("dtype", [np.uint16, np.float32])
@pytest.mark.parametrize("f", [FilterType.Triangle, FilterType.CatmullRom, FilterType.Gaussian, FilterType.Lanczos3])
@pytest.mark.parametrize("r", [2, 4])
def test_resize_integer_ratio_rgba16_rgba32f(cuda, f, r, dtype):
    """The four-channel fixed-window downscale kernel with 16-bit and f32 pixels (same passes, other load and store): batches
    (so that it is picked), ragged strips and bands, both accumulator start values; +-1 LSB resp. 1e-5 relative."""
    import torch
    rng = np.random.default_rng(700 + 10 * r + int(f))
    lib = L.lib()
    st = torch.cuda.current_stream().cuda_stream
    dt = L.U16 if dtype == np.uint16 else L.F32
    for (nw, nh, n) in [(372, 92, 12), (960, 540, 1), (50, 311, 40)]:
        w, h = nw * r, nh * r
        imgs = np.stack([rand_image(rng, h, w, 4, dtype) for _ in range(n)])
        src = torch.from_numpy(imgs).to(cuda)
        for acc in (1.0, 0.0):
            dst = torch.empty((n, nh, nw, 4), dtype=src.dtype, device=cuda)
            L.check(lib.b200img_resize_batch_dev(src.data_ptr(), 0, w, h, 4, dt, dst.data_ptr(), 0, nw, nh, n, int(f), acc, L.FP_FAST, st))
            got = dst.cpu().numpy()
            for i in (0, n - 1):
                want = o.resize(imgs[i], nw, nh, int(f), acc_init=acc)
                if dtype == np.float32:
                    assert np.allclose(got[i], want, rtol=1e-5, atol=1e-6), (f.name, r, nw, nh, acc)
                else:
                    assert np.abs(got[i].astype(np.int64) - want.astype(np.int64)).max() <= 1, (f.name, r, nw, nh, acc)
    if dtype == np.float32:   # non-finite samples: every window holds exactly the reference's taps, the padding is finite zeros
        imgs = rng.random((12, 92 * r, 372 * r, 4), dtype=np.float32) * 0.5 + 0.1
        imgs[0, 40 * r, 100 * r] = np.inf
        imgs[0, 2, 3, 1] = np.nan
        imgs[11, 92 * r - 1, 372 * r - 1, 2] = np.nan
        src = torch.from_numpy(imgs).to(cuda)
        dst = torch.empty((12, 92, 372, 4), dtype=torch.float32, device=cuda)
        L.check(lib.b200img_resize_batch_dev(src.data_ptr(), 0, 372 * r, 92 * r, 4, L.F32, dst.data_ptr(), 0, 372, 92, 12, int(f), 0.0, L.FP_FAST, st))
        got = dst.cpu().numpy()
        for i in (0, 11):
            want = o.resize(imgs[i], 372, 92, int(f), acc_init=0.0)
            assert np.array_equal(got[i] == 0, want == 0)
            assert np.allclose(got[i], want, rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("dtype", [np.uint8, np.uint16, np.float32])
@pytest.mark.parametrize("f", [FilterType.Triangle, FilterType.CatmullRom, FilterType.Gaussian, FilterType.Lanczos3])
def test_resize_ratio3_four_channels(cuda, f, dtype):
    """Downscale by 3 (an odd ratio centres every window on a sample: one tap more, one more open output on the rows that close
    one) through the four-channel fixed-window kernel: batches, ragged strips and bands, both accumulator start values."""
    import torch
    rng = np.random.default_rng(900 + int(f))
    lib = L.lib()
    st = torch.cuda.current_stream().cuda_stream
    dt = {np.uint8: L.U8, np.uint16: L.U16, np.float32: L.F32}[dtype]
    for (nw, nh, n) in [(373, 91, 12), (640, 360, 1), (47, 311, 40), (1280, 720, 2)]:
        w, h = nw * 3, nh * 3
        imgs = np.stack([rand_image(rng, h, w, 4, dtype) for _ in range(n)])
        src = torch.from_numpy(imgs).to(cuda)
        for acc in (1.0, 0.0):
            dst = torch.empty((n, nh, nw, 4), dtype=src.dtype, device=cuda)
            L.check(lib.b200img_resize_batch_dev(src.data_ptr(), 0, w, h, 4, dt, dst.data_ptr(), 0, nw, nh, n, int(f), acc, L.FP_FAST, st))
            general = torch.empty_like(dst)
            L.check(lib.b200img_resize_batch_dev(src.data_ptr(), 0, w, h, 4, dt, general.data_ptr(), 0, nw, nh, n, int(f), acc,
                                                 L.FP_FAST | L.RESIZE_GENERAL, st))
            got, gen = dst.cpu().numpy(), general.cpu().numpy()
            for i in (0, n - 1):
                want = o.resize(imgs[i], nw, nh, int(f), acc_init=acc)
                if dtype == np.float32:
                    assert np.allclose(got[i], want, rtol=1e-5, atol=1e-6), (f.name, nw, nh, acc)
                    assert np.allclose(gen[i], want, rtol=1e-5, atol=1e-6)
                else:
                    assert np.abs(got[i].astype(np.int64) - want.astype(np.int64)).max() <= 1, (f.name, nw, nh, acc)
                    assert np.abs(gen[i].astype(np.int64) - want.astype(np.int64)).max() <= 1
