"""Pins the CPU oracle against every known-answer test the reference holds for the hot path
(SURVEY.md section 8c).  CPU only."""
import numpy as np
import pytest

import oracle as o
from conftest import Lcg

U8, U16, F32 = np.uint8, np.uint16, np.float32


# ---- conversions: src/metadata/cicp.rs:1328-1390 ----
def test_cicp_transform_pixels_srgb():
    data = np.array([255, 0, 0, 255], U8)
    assert o.convert(data, 4, U8, 3).tolist() == [[255, 0, 0]]
    assert o.convert(data, 4, U8, 1).tolist() == [[54]]  # 255 * 0.2126
    assert o.convert(data, 4, U8, 2).tolist() == [[54, 255]]


@pytest.mark.parametrize("v,exp", [(65535 // 2, 127), (65535 // 2 + 1, 128)])
def test_cicp_transform_pixels_srgb_16(v, exp):
    data = np.array([v], U16)
    assert o.convert(data, 1, U8, 3).tolist() == [[exp] * 3]
    assert o.convert(data, 1, U8, 1).tolist() == [[exp]]
    assert o.convert(data, 1, U8, 2).tolist() == [[exp, 255]]
    la = np.array([v, 65535], U16)
    assert o.convert(la, 2, U8, 3).tolist() == [[exp] * 3]
    assert o.convert(la, 2, U8, 1).tolist() == [[exp]]
    assert o.convert(la, 2, U8, 2).tolist() == [[exp, 255]]


# src/images/dynimage.rs:2187-2203, 2310-2327
def test_dynimage_fastpath_and_luma():
    assert o.convert(np.array([65535, 0, 0], U16), 3, U8, 4).tolist() == [[255, 0, 0, 255]]
    assert o.convert(np.array([65535, 0, 0], U16), 3, U8, 1).tolist() == [[54]]
    p3 = [0.22897343, 0.69173974, 0.07928687]  # Display P3 luminance row (D65)
    assert o.convert(np.array([65535, 0, 0], U16), 3, U8, 1, coefs=p3).tolist() == [[58]]
    assert o.convert(np.array([65535], U16), 1, U8, 3).tolist() == [[255, 255, 255]]


# src/images/dynimage.rs:2384-2418 (sRGB / Unspecified rows; other primaries need moxcms)
def test_from_rgb_srgb_luma_table():
    assert o.convert(np.array([54, 182, 18], U8), 3, U8, 1).tolist() == [[143]]
    assert o.convert(np.array([0, 255, 0], U8), 3, U8, 1, coefs=[0.0, 1.0, 0.0]).tolist() == [[255]]


# src/color.rs:1123-1151
@pytest.mark.parametrize("c", [1, 2, 3, 4])
def test_lossless_u8_u16_u8(c):
    a = np.full((1, c), 255 >> 2, U8)
    b = o.convert(a, c, U16, c, o.CONVERT)
    assert o.convert(b, c, U8, c, o.CONVERT).tolist() == a.tolist()


def test_accuracy_conversion():
    assert o.convert(np.array([13, 13, 13], U8), 3, U8, 1, o.CONVERT).tolist() == [[13]]


# tests/conversions.rs:10-40
@pytest.mark.parametrize("c", [3, 4])
def test_u8_max_to_u16_max(c):
    a = np.full((4, c), 255, U8)
    assert (o.convert(a, c, U16, c, o.CONVERT) == 65535).all()


def test_from_primitive_rules():
    # src/color.rs:540-622
    assert o.convert(np.array([np.nan, -1.0, 2.0, 0.5], F32), 1, U8, 1).ravel().tolist() == [255, 0, 255, 128]
    assert o.convert(np.array([np.nan, -1.0, 2.0], F32), 1, U16, 1).ravel().tolist() == [65535, 0, 65535]
    u16 = np.arange(65536, dtype=U16)
    assert np.array_equal(o.convert(u16, 1, U8, 1).ravel(), ((u16.astype(np.uint32) + 128) // 257).astype(U8))
    assert np.array_equal(o.convert(u16, 1, F32, 1).ravel(), u16.astype(F32) / F32(65535))
    u8 = np.arange(256, dtype=U8)
    assert np.array_equal(o.convert(u8, 1, F32, 1).ravel(), u8.astype(F32) / F32(255))
    assert np.array_equal(o.convert(u8, 1, U16, 1).ravel(), u8.astype(U16) * 257)


def test_u8_luma_matches_its_documented_formula():
    # src/primitive_sealed.rs:203-212: "same results as (r*2126 + g*7152 + b*722 + 5000) / 10000"
    rng = np.random.default_rng(1)
    rgb = rng.integers(0, 256, (20000, 3), dtype=U8)
    got = o.convert(rgb, 3, U8, 1, o.CONVERT).ravel()
    r, g, b = (rgb[:, i].astype(np.uint32) for i in range(3))
    assert np.array_equal(got, ((r * 2126 + g * 7152 + b * 722 + 5000) // 10000).astype(U8))


# ---- fast_blur: src/imageops/fast_blur.rs:480-507 (division identity) ----
def test_u8_weight_exhaustive_small():
    for ks in range(3, 52, 2):
        for acc in range(0, ks * 255 + 1, 7):
            assert o.u8_box_store(acc, ks) == ((acc + ks // 2) // ks) & 0xFF
        assert o.u8_box_store(ks * 255, ks) == 255


def test_u8_weight_sampled_large():
    for ks in range(53, 1026, 36):
        mx = ks * 255
        for acc in range(0, mx + 1, max(mx // 500, 1)):
            assert o.u8_box_store(acc, ks) == ((acc + ks // 2) // ks) & 0xFF


# fast_blur.rs:535-592: a constant image stays constant for sigma in (0, 100)
def test_box_blur_constant_image():
    g = Lcg(0xC0FFEE)
    done = 0
    while done < 20:
        w, h = g.next_u32() % 256, g.next_u32() % 256
        sigma = 100.0 * g.next_u32() / 2.0 ** 32
        px, cn = g.next_u32() % 256, g.next_u32() % 4 + 1
        if w == 0 or h == 0 or sigma <= 0:
            continue
        img = np.full((h, w, cn), px, U8)
        assert (o.fast_blur(img, sigma) == px).all()
        done += 1


def test_boxes_for_gauss():
    assert o.boxes_for_gauss(10.0) == [9, 9, 11]
    assert o.boxes_for_gauss(0.0) == [1, 1, 1]


# src/imageops/mod.rs:473-555
def test_fast_blur_edge_cases():
    img = np.zeros((7, 5, 3), U8)
    for s in (0.0, -1.0, 50.0):
        assert o.fast_blur(img, s).shape == img.shape
    assert o.fast_blur(np.zeros((0, 0, 3), U8), 3.0).size == 0


def test_fast_blur_approximates_gaussian():
    rng = np.random.default_rng(3)
    base = rng.integers(0, 65536, (8, 8, 3), dtype=U16)
    img = np.kron(base, np.ones((16, 16, 1), U16))  # 128x128 blocky image
    a = o.fast_blur(img, 6.0).astype(np.float64)
    b = o.gaussian_blur(img, o.gauss_params_from_sigma(6.0)).astype(np.float64)
    assert (np.abs(a - b) / np.maximum(b, 1)).mean() < 0.05


# ---- Gaussian parameters: src/imageops/sample.rs:1132-1294 ----
def test_gauss_params():
    p = o.gauss_params_from_radius(10.0)
    assert (p.x_kernel_size, p.y_kernel_size) == (21, 21) and abs(p.x_sigma - 3.5) < 1e-6
    p = o.gauss_params_from_sigma(10.0)
    assert p.x_kernel_size == 65 and p.x_sigma == 10.0
    p = o.gauss_params_from_sigma(0.1)
    assert (p.x_kernel_size, p.x_sigma) == (1, 1.0)
    p = o.gauss_params_from_radius(0.0)
    assert p.x_kernel_size == 1 and abs(p.x_sigma - (0.3 * ((0.8 - 1) * 0.5 - 1) + 0.8)) < 1e-6
    for bad in (-1.0, float("nan"), float("inf")):
        with pytest.raises(ValueError):
            o.gauss_params_from_radius(bad)
        with pytest.raises(ValueError):
            o.gauss_params_from_sigma(bad)
    k = o.gaussian_kernel_1d(21, 3.5)
    assert abs(k.sum() - 1) < 1e-6 and np.allclose(k, k[::-1]) and k.argmax() == 10


def test_gaussian_constant_and_identity():
    img = np.full((9, 11, 4), 201, U8)
    for path in ("generic", "dyn"):
        assert (o.gaussian_blur(img, o.gauss_params_from_sigma(2.5), path) == 201).all()
    # 1-tap quirks (SURVEY hard part 8): row kernel shifts right by one, ring-queue column leaves row 0 zero
    rng = np.random.default_rng(5)
    im = rng.integers(0, 256, (6, 7, 1), dtype=U8)
    out = o.gaussian_blur(im, o.gauss_params_from_radius(0.0), "generic")
    assert (out[0] == 0).all()
    assert np.array_equal(out[1:, 1:], im[1:, :-1]) and np.array_equal(out[1:, 0], im[1:, 0])


# ---- resampler invariants: src/imageops/sample.rs:1812-1919 ----
def test_issue_186_and_bug_1600():
    assert o.resize(np.zeros((100, 100, 3), U8), 50, 50, "lanczos3").shape == (50, 50, 3)
    big = np.full((627, 629, 4), 255, U8)
    assert o.resize(big, 22, 22, "lanczos3").any()


def test_issue_2340_empty_sources():
    for shape in ((0, 1 << 20, 1), (1 << 20, 0, 1)):
        empty = np.zeros(shape, U8)
        assert not o.resize(empty, 1, 1, "lanczos3").any()
        assert o.resize(empty, 256, 256, "lanczos3").shape == (256, 256, 1)


def test_resize_same_size_is_copy():
    rng = np.random.default_rng(0)
    img = rng.integers(0, 256, (13, 17, 3), dtype=U8)
    assert np.array_equal(o.resize(img, 17, 13, "catmullrom"), img)


def test_resize_transparent_image_invariant():
    # sample.rs:1861-1897 on a synthetic stand-in for tp1n3p08.png: opaque centre, transparent rim
    img = np.zeros((32, 32, 4), U8)
    img[4:28, 4:28] = [200, 100, 50, 255]
    for f in ("nearest", "triangle", "catmullrom", "gaussian", "lanczos3"):
        a = o.resize(img, 16, 16, f)[5:11, 5:11, 3]
        assert not np.isin(a, [253, 254]).any(), f


def test_acc_init_finding_f2():
    # the literal reference adds 1 (accumulator starts at 1.0)
    black = np.zeros((20, 20, 3), U8)
    assert (o.resize(black, 10, 10, "lanczos3") == 1).all()
    assert (o.resize(black, 10, 10, "lanczos3", acc_init=0.0) == 0).all()
    f = np.full((8, 8, 4), 0.25, F32)
    assert (o.resize(f, 4, 4, "triangle") == 1.0).all()  # clamp_nearest_from(1.25)
    assert np.allclose(o.resize(f, 4, 4, "triangle", acc_init=0.0), 0.25)


def test_nearest_roundtrip_identity():
    # tests/regression.rs:154-174 (2x up then down == identity) holds with acc_init = 0
    rng = np.random.default_rng(2)
    img = rng.integers(0, 256, (33, 65, 4), dtype=U8)
    up = o.resize(img, 130, 66, "nearest", acc_init=0.0)
    assert np.array_equal(o.resize(up, 65, 33, "nearest", acc_init=0.0), img)


def test_filter_kernels():
    assert o.filter_kernel("lanczos3", 0.0) == 1.0 and o.filter_kernel("lanczos3", 3.0) == 0.0
    assert o.filter_kernel("catmullrom", 0.0) == 1.0 and o.filter_kernel("catmullrom", 1.0) == 0.0
    assert o.filter_kernel("triangle", 0.25) == 0.75 and o.filter_kernel("nearest", 9.0) == 1.0
    assert [o.filter_support(f) for f in range(5)] == [0.0, 1.0, 2.0, 3.0, 3.0]
    l, c, w = o.resize_weights("nearest", 10, 4, 0)
    assert (c == 1).all() and (w[:, 0] == 1.0).all() and l.tolist() == [1, 3, 6, 8]


# src/math/utils.rs:86-112 (quickcheck properties) + resize_handles_fill
def test_resize_dimensions():
    rng = np.random.default_rng(11)
    for _ in range(300):
        old_w, new_w = int(rng.integers(1, 2 ** 20)), int(rng.integers(1, 2 ** 20))
        if new_w * 400 >= old_w * (2 ** 32 - 1):
            continue
        rw, rh = o.resize_dimensions(old_w, 400, new_w, 2 ** 32 - 1, False)
        exact = max(int(np.floor(400.0 * new_w / old_w + 0.5)), 1)
        assert rw == new_w and rh == exact
    assert o.resize_dimensions(100, 200, 200, 500, True) == (250, 500)
    assert o.resize_dimensions(200, 100, 500, 200, True) == (500, 250)
    assert o.resize_dimensions(4032, 3024, 400, 400, False) == (400, 300)


def test_swizzle():
    a = np.arange(14, dtype=U8)
    assert o.swizzle_02(a, 4).tolist() == [2, 1, 0, 3, 6, 5, 4, 7, 10, 9, 8, 11, 12, 13]
    assert o.swizzle_02(a, 3).tolist() == [2, 1, 0, 5, 4, 3, 8, 7, 6, 11, 10, 9, 12, 13]


# ---- filter3x3 / unsharpen restatements (sample.rs:897-956, 1617-1654) -------------------
def test_filter3x3_oracle_invariants():
    rng = np.random.default_rng(5)
    img = rng.integers(0, 256, (9, 11, 3), dtype=np.uint8)
    ident = o.filter3x3(img, [0, 0, 0, 0, 1, 0, 0, 0, 0])
    assert np.array_equal(ident[1:-1, 1:-1], img[1:-1, 1:-1])
    assert not ident[0].any() and not ident[-1].any() and not ident[:, 0].any() and not ident[:, -1].any()
    # weights are normalised by their sum: a box kernel on a constant image gives the constant back
    const = np.full((6, 7, 1), 77, np.uint8)
    assert (o.filter3x3(const, [1] * 9)[1:-1, 1:-1] == 77).all()
    # a zero-sum kernel is not normalised (inverse_sum = 1) and negative sums clamp to 0
    edge = o.filter3x3(const, [1, 1, 1, 1, -8, 1, 1, 1, 1])
    assert not edge.any()
    # plain numpy restatement on f32, same tap order
    f = rng.random((8, 9, 2), dtype=np.float32)
    k = np.array([0.5, 1, 0.25, 2, 3, 1, 0.125, 1, 0.5], np.float32)
    t = np.zeros((6, 7, 2), np.float32)
    for j in range(3):
        for i in range(3):
            t = t + f[j:j + 6, i:i + 7] * k[j * 3 + i]
    s = np.float32(0)
    for v in k:
        s = np.float32(s + v)
    want = np.clip(t * (np.float32(1) / s), 0, 1).astype(np.float32)
    assert np.array_equal(o.filter3x3(f, k)[1:-1, 1:-1], want)


def test_unsharpen_oracle_invariants():
    rng = np.random.default_rng(6)
    const = np.full((12, 10, 3), 90, np.uint8)
    assert np.array_equal(o.unsharpen(const, 2.0, 0), const)
    img = rng.integers(0, 256, (14, 15, 3), dtype=np.uint8)
    # a threshold nothing can exceed returns the image itself
    assert np.array_equal(o.unsharpen(img, 1.5, 255), img)
    blurred = o.gaussian_blur(img, o.gauss_params_from_sigma(1.5))
    d = img.astype(np.int32) - blurred.astype(np.int32)
    want = np.where(np.abs(d) > 3, np.clip(img.astype(np.int32) + d, 0, 255), img).astype(np.uint8)
    assert np.array_equal(o.unsharpen(img, 1.5, 3), want)


# ---- thumbnail restatement (sample.rs:564-886) -------------------------------------------
def test_thumbnail_reference_kat_huge_block():
    """sample.rs:1922-1930 `thumbnail_huge_block`: neither input may overflow or panic."""
    assert o.thumbnail(np.full((4113, 4096, 1), 255, np.uint8), 1, 1)[0, 0, 0] == 255
    assert o.thumbnail(np.full((1024, 1024, 1), 65535, np.uint16), 1, 1)[0, 0, 0] == 65535


def test_thumbnail_oracle_invariants():
    rng = np.random.default_rng(7)
    img = rng.integers(0, 256, (40, 60, 3), dtype=np.uint8)
    want = (img.reshape(20, 2, 30, 2, 3).astype(np.uint32).sum(axis=(1, 3)) + 2) // 4
    assert np.array_equal(o.thumbnail(img, 30, 20), want)
    assert np.array_equal(o.thumbnail(img, 60, 40), img)          # same size: 1x1 blocks
    const = np.full((10, 10, 3), 77, np.uint8)
    for size in ((20, 20), (13, 7), (3, 25), (10, 10)):           # all four window forms
        assert (o.thumbnail(const, *size) == 77).all()
    assert not o.thumbnail(np.zeros((0, 4, 1), np.uint8), 3, 3).any()
    with pytest.raises(o.ReferencePanic):                         # integer division by a zero count
        o.thumbnail(np.zeros((1000, 100, 1), np.uint8), 1, 1)


# ---- colorops restatements, pinned by colorops.rs:657-700 --------------------------------
def test_colorops_reference_kats():
    img = np.array([0, 1, 2, 10, 11, 12], np.uint8).reshape(2, 3, 1)
    assert o.grayscale(img).ravel().tolist() == [0, 1, 2, 10, 11, 12]                 # test_grayscale
    assert o.invert(img).ravel().tolist() == [255, 254, 253, 245, 244, 243]           # test_invert
    assert o.brighten(img, 10).ravel().tolist() == [10, 11, 12, 20, 21, 22]           # test_brighten(_place)


def test_colorops_oracle_invariants():
    rng = np.random.default_rng(8)
    rgba = rng.integers(0, 256, (5, 6, 4), dtype=np.uint8)
    assert np.array_equal(o.invert(rgba)[..., 3], rgba[..., 3])                       # alpha kept
    assert np.array_equal(o.invert(o.invert(rgba)), rgba)
    assert np.array_equal(o.brighten(rgba, 300)[..., :3], np.full((5, 6, 3), 255))
    assert np.array_equal(o.huerotate(rgba, 0), rgba)
    la = rng.integers(0, 65536, (4, 4, 2), dtype=np.uint16)
    assert np.array_equal(o.huerotate(la, 90), la)                                    # < 3 channels: copy
    rgb16 = rng.integers(0, 65536, (4, 4, 3), dtype=np.uint16)
    assert o.huerotate(rgb16, 30).max() <= 255                                        # the literal clamp to 255
    f = np.array([[[0.2, 0.7, 1.0, 0.5]]], np.float32)
    assert o.brighten(f, 1).ravel().tolist() == [1.0, 1.0, 1.0, 0.5]                  # f32 goes through the i32 cast
    m = o.huerotate_matrix(0)
    assert np.allclose(m.reshape(3, 3), np.eye(3), atol=1e-12)


def test_has_constant_alpha_oracle():
    """resize.rs:155-218"""
    rgba = np.zeros((6, 7, 4), np.uint8)
    rgba[..., 3] = 200
    assert o.has_constant_alpha(rgba)
    rgba[5, 6, 3] = 201
    assert not o.has_constant_alpha(rgba)
    assert o.has_constant_alpha(np.arange(60, dtype=np.uint16).reshape(4, 5, 3))      # no alpha channel
    assert o.has_constant_alpha(np.zeros((0, 3, 2), np.float32))                      # empty
    nan = np.full((2, 2, 2), np.nan, np.float32)
    assert not o.has_constant_alpha(nan)


def test_chunked_filters_and_region_resampler_equal_the_checker():
    """The timed CPU baseline (filters with the reference's 4- / 16- / 32-wide chunk structure, vectorised by the
    compiler) and the windowed resampler used for full-size spot checks are bit-identical to the scalar checker."""
    rng = np.random.default_rng(77)
    for dt in (np.uint8, np.uint16, np.float32):
        img = rng.random((37, 53, 3), dtype=np.float32) if dt == np.float32 else rng.integers(0, np.iinfo(dt).max + 1, (37, 53, 3), dtype=dt)
        for sigma in (0.1, 1.3, 9.0):
            p = o.gauss_params_from_sigma(sigma)
            for path in ("generic", "dyn"):
                assert o.gaussian_blur(img, p, path).tobytes() == o.gaussian_blur(img, p, path, chunked=True).tobytes()
        for f in ("lanczos3", "catmullrom", "nearest"):
            full = o.resize(img, 91, 23, f)
            reg = o.resize_region(lambda r0, r1, c0, c1: img[r0:r1, c0:c1], 53, 37, 3, dt, 91, 23, f, 2, 21, 5, 90, 1.0)
            assert reg.tobytes() == full[2:21, 5:90].tobytes()
