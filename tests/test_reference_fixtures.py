"""The reference's OWN test images through the oracle (CPU) and the CUDA path (GPU).

tests/golden/reference_fixtures.npz holds the decoded pixels of the fixture files the reference's tests open
(made by tests/golden/make_reference_fixtures.py where /root/reference exists; the GPU box has no reference tree).
Each test below restates one reference test on the real file:

  resizing_with_alpha                        tests/regression.rs:154-174
  resizing_with_catmul (informational)       tests/regression.rs:188-214 -- DynamicImage::resize_exact runs the
                                             unvendored pic-scale-safe crate there, so its golden PNG cannot gate the
                                             in-tree resampler; the difference is measured and bounded, not asserted equal
  resize_transparent_image                   src/imageops/sample.rs:1861-1897
  fast_blur_approximates_gaussian_blur_well  src/imageops/mod.rs:535-555
  examples/scaleup (tinycross.png)           examples/scaleup/main.rs
"""
import json
import os

import numpy as np
import pytest

import oracle as o

HERE = os.path.dirname(os.path.abspath(__file__))
FILTERS = ["nearest", "triangle", "catmullrom", "gaussian", "lanczos3"]


@pytest.fixture(scope="module")
def fx():
    return dict(np.load(os.path.join(HERE, "golden", "reference_fixtures.npz")))


def test_fixture_file_matches_the_reference_tree(fx):
    """Where the reference tree is present (this container), the committed pixels are the files' pixels."""
    ref = "/root/reference"
    if not os.path.isdir(ref):
        pytest.skip("no reference tree here")
    from PIL import Image
    a = np.asarray(Image.open(os.path.join(ref, "tests/regression/image/resize-with-alpha-original.png")))
    assert np.array_equal(a, fx["alpha_original"])
    a = np.asarray(Image.open(os.path.join(ref, "tests/images/png/transparency/tp1n3p08.png")).convert("RGBA"))
    assert np.array_equal(a, fx["tp1n3p08"])
    assert fx["rgb16"].shape == (151, 157, 3) and fx["rgb16"].dtype == np.uint16 and fx["rgb16"].max() > 255


# ---- the four reference tests as functions of a `resize / blur / fast_blur / to_rgb8` implementation ----
def check_resizing_with_alpha(img, resize):
    h, w = img.shape[:2]
    up = resize(img, 2 * w, 2 * h, "nearest", 0.0)
    back = resize(up, w, h, "nearest", 0.0)
    assert np.array_equal(back, img)   # the identity the reference asserts (accumulator starting at 0)
    # the in-tree resampler's literal accumulator (1.0, sample.rs:547) adds one level per pass, saturating
    lit = resize(resize(img, 2 * w, 2 * h, "nearest", 1.0), w, h, "nearest", 1.0)
    assert np.array_equal(lit, np.minimum(img.astype(int) + 2, 255).astype(np.uint8))


def check_resize_transparent_image(img, resize):
    for f in FILTERS:
        out = resize(img, 16, 16, f, 1.0)   # imageops::resize itself: the literal arithmetic
        alpha = out[5:11, 5:11, 3]
        assert not np.isin(alpha, (253, 254)).any(), (f, alpha)


def catmul_difference(img, golden, resize):
    out = {}
    for acc in (0.0, 1.0):
        got = resize(img, golden.shape[1], golden.shape[0], "catmullrom", acc)
        d = np.abs(got.astype(int) - golden.astype(int))
        opaque = golden[:, :, 3] == 255
        out[f"acc_init_{acc:g}"] = {"max_abs": int(d.max()), "exact": float((d == 0).mean()),
                                    "max_abs_opaque_pixels": int(d[opaque].max()), "exact_opaque_pixels": float((d[opaque] == 0).mean())}
    return out


def check_fast_blur_approximates_gaussian(img16, blur_dyn_sigma, fast_blur, to_rgb8):
    gauss = to_rgb8(blur_dyn_sigma(img16, 50.0)).astype(np.float32).reshape(-1)
    fast = to_rgb8(fast_blur(img16, 50.0)).astype(np.float32).reshape(-1)
    with np.errstate(divide="ignore", invalid="ignore"):
        err = np.float32(((gauss - fast) / gauss).sum(dtype=np.float32)) / np.float32(gauss.size)
    assert err < 0.05, err
    return float(err)


# ---- oracle (CPU) ----
def _o_resize(img, nw, nh, f, acc):
    return o.resize(img, nw, nh, f, acc_init=acc)


def test_oracle_resizing_with_alpha(fx):
    check_resizing_with_alpha(fx["alpha_original"], _o_resize)


def test_oracle_resize_transparent_image(fx):
    check_resize_transparent_image(fx["tp1n3p08"], _o_resize)


def test_oracle_catmul_difference_is_bounded(fx):
    d = catmul_difference(fx["alpha_original"], fx["alpha_half"], _o_resize)
    print("resizing_with_catmul, in-tree resampler vs the pic-scale-safe golden:", json.dumps(d))
    # informational, with a loose floor so that a broken resampler still shows: opaque pixels do not depend on
    # the premultiplication pic-scale-safe adds, and there the intended filter (accumulator from 0) reproduces
    # > 80 % of the golden's bytes exactly (the rest differ by pic-scale-safe's own fixed-point weights) while
    # the literal accumulator (from 1.0) reproduces about half as many
    assert d["acc_init_0"]["exact_opaque_pixels"] > 0.8 > d["acc_init_1"]["exact_opaque_pixels"]


def test_oracle_fast_blur_approximates_gaussian(fx):
    check_fast_blur_approximates_gaussian(
        fx["rgb16"], lambda a, s: o.gaussian_blur(a, o.gauss_params_from_sigma(s), "dyn"), o.fast_blur,
        lambda a: o.convert(a.reshape(-1, 3), 3, np.uint8, 3).reshape(a.shape))


def test_scaledown_report_is_committed():
    r = json.load(open(os.path.join(HERE, "golden", "reference_fixture_report.json")))
    for k in ("near", "tri", "cmr", "gauss", "lcz2"):
        assert r[k]["acc_init_0"]["within_2"] > 0.98   # informational numbers, see the generating script


# ---- CUDA path (GPU), through the C ABI ----
def _g_resize_factory(flags):
    from image_b200 import FilterType, imageops
    ft = {"nearest": FilterType.Nearest, "triangle": FilterType.Triangle, "catmullrom": FilterType.CatmullRom,
          "gaussian": FilterType.Gaussian, "lanczos3": FilterType.Lanczos3}
    return lambda img, nw, nh, f, acc: imageops.resize(np.ascontiguousarray(img), nw, nh, ft[f], acc_init=acc, flags=flags)


@pytest.mark.gpu
@pytest.mark.parametrize("strict", [True, False])
def test_gpu_reference_resize_tests(cuda, fx, strict):
    from image_b200 import _lib as L
    rs = _g_resize_factory(L.FP_STRICT if strict else L.FP_FAST)
    check_resizing_with_alpha(fx["alpha_original"], rs)
    check_resize_transparent_image(fx["tp1n3p08"], rs)
    d = catmul_difference(fx["alpha_original"], fx["alpha_half"], rs)
    if strict:
        assert d == catmul_difference(fx["alpha_original"], fx["alpha_half"], _o_resize)
    # every filter on the real files, against the oracle: bit-exact (STRICT) / +-1 (FAST)
    for name, sizes in (("alpha_original", [(65, 33), (258, 132)]), ("tp1n3p08", [(16, 16), (48, 48)]),
                        ("tinycross", [(32, 32), (48, 48)]), ("fractal", [(400, 400), (1200, 900)])):
        for nw, nh in sizes:
            for f in FILTERS:
                want = o.resize(fx[name], nw, nh, f)
                got = rs(fx[name], nw, nh, f, 1.0)
                if strict:
                    assert np.array_equal(got, want), (name, f, nw, nh)
                else:
                    assert np.abs(got.astype(int) - want.astype(int)).max() <= 1, (name, f, nw, nh)


@pytest.mark.gpu
def test_gpu_fast_blur_approximates_gaussian(cuda, fx):
    from image_b200 import GaussianBlurParameters, imageops
    err = check_fast_blur_approximates_gaussian(
        fx["rgb16"], lambda a, s: imageops.blur_dyn(a, GaussianBlurParameters.new_from_sigma(s)), imageops.fast_blur,
        lambda a: imageops.convert_pixels(a, 3, np.uint8))
    # and the same three images as the oracle, bit for bit (Q0.15 blur, u16 box blur, u16 -> u8 are all integer paths)
    img = fx["rgb16"]
    assert np.array_equal(imageops.blur_dyn(img, GaussianBlurParameters.new_from_sigma(50.0)),
                          o.gaussian_blur(img, o.gauss_params_from_sigma(50.0), "dyn"))
    assert np.array_equal(imageops.fast_blur(img, 50.0), o.fast_blur(img, 50.0))
    assert err < 0.05


@pytest.mark.gpu
def test_gpu_blurs_on_a_real_image(cuda, fx):
    from image_b200 import GaussianBlurParameters, imageops
    from image_b200 import _lib as L
    img = fx["fractal"]
    for sigma in (1.0, 3.5, 10.0):
        p, op = GaussianBlurParameters.new_from_sigma(sigma), o.gauss_params_from_sigma(sigma)
        assert np.array_equal(imageops.blur_dyn(img, p), o.gaussian_blur(img, op, "dyn"))
        want = o.gaussian_blur(img, op, "generic")
        assert np.array_equal(imageops.blur_advanced(img, p, flags=L.FP_STRICT), want)
        d = np.abs(imageops.blur_advanced(img, p).astype(int) - want.astype(int))
        assert d.max() <= 1
        assert np.array_equal(imageops.fast_blur(img, sigma), o.fast_blur(img, sigma))
