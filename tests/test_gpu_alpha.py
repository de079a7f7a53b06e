"""GPU parity for the steps either side of DynamicImage::resize."""
import numpy as np
import pytest

import oracle as o
from conftest import rand_image, same
from image_b200 import FilterType, imageops
from image_b200.images import DynamicImage, ImageBuffer

pytestmark = pytest.mark.gpu
DT = [np.uint8, np.uint16, np.float32]


@pytest.mark.parametrize("dtype", DT)
@pytest.mark.parametrize("c", [1, 2, 3, 4])
def test_has_constant_alpha(cuda, dtype, c):
    import torch
    rng = np.random.default_rng(120 + c)
    for (h, w) in ((31, 45), (64, 64), (1, 1), (3, 1000)):
        img = rand_image(rng, h, w, c, dtype)
        const = img.copy()
        if c in (2, 4):
            const[..., c - 1] = const[0, 0, c - 1]
        one_off = const.copy()
        if c in (2, 4):
            one_off[h - 1, w - 1, c - 1] = one_off[0, 0, c - 1] + (1 if dtype != np.float32 else 0.25)
        for a in (img, const, one_off):
            want = o.has_constant_alpha(a)
            assert imageops.has_constant_alpha(a) == want
            assert imageops.has_constant_alpha(torch.from_numpy(a).to(cuda)) == want
    assert imageops.has_constant_alpha(np.zeros((0, 4, 4), dtype)) is True      # empty image
    if dtype == np.float32 and c in (2, 4):
        nan = np.zeros((4, 4, c), np.float32)
        nan[..., c - 1] = np.nan                                                # NaN == NaN is false (resize.rs:214-217)
        assert o.has_constant_alpha(nan) is False and imageops.has_constant_alpha(nan) is False
        zeros = np.zeros((4, 4, c), np.float32)
        zeros[1, 1, c - 1] = -0.0                                               # -0.0 == 0.0
        assert imageops.has_constant_alpha(zeros) is True


@pytest.mark.parametrize("dtype", DT)
@pytest.mark.parametrize("c", [1, 2, 3, 4])
def test_premultiply_roundtrip_against_own_rule(cuda, dtype, c):
    """pic-scale-safe is unvendored: this pins the documented rule, not the reference"""
    import torch
    rng = np.random.default_rng(130 + c)
    img = rand_image(rng, 29, 37, c, dtype)
    if c in (2, 4):
        img[0, :5, c - 1] = 0        # a = 0 rows
    pre = o.alpha_mul(img, False)
    dev = torch.from_numpy(img.copy()).to(cuda)
    imageops.premultiply_alpha(dev)
    assert same(dev.cpu().numpy(), pre)
    host = img.copy()
    imageops.premultiply_alpha(host)
    assert same(host, pre)
    imageops.unpremultiply_alpha(dev)
    assert same(dev.cpu().numpy(), o.alpha_mul(pre, True))
    if c in (1, 3):
        assert same(pre, img)       # resize.rs:105-129: Luma / Rgb are left alone


def test_dynamic_image_resize_control_flow(cuda):
    """resize.rs:13-93: same size -> unchanged; Nearest or constant alpha -> no premultiply"""
    rng = np.random.default_rng(140)
    img = rand_image(rng, 40, 60, 4, np.uint8)
    d = DynamicImage(ImageBuffer(img))
    assert d.resize_exact(60, 40, FilterType.Lanczos3) is d
    near = d.resize_exact(30, 20, FilterType.Nearest).buf.data
    assert same(near, o.resize(img, 30, 20, "nearest"))
    opaque = img.copy()
    opaque[..., 3] = 255
    got = DynamicImage(ImageBuffer(opaque)).resize_exact(30, 20, FilterType.CatmullRom).buf.data
    assert np.abs(got.astype(int) - o.resize(opaque, 30, 20, "catmullrom").astype(int)).max() <= 1
    # varying alpha: premultiply -> sample.rs resampler -> unpremultiply, composed from the oracle's pieces
    want = o.alpha_mul(o.resize(o.alpha_mul(img, False), 30, 20, "catmullrom"), True)
    got = d.resize_exact(30, 20, FilterType.CatmullRom).buf.data
    # the +-1 LSB of the FAST resampler is amplified by max/a in the unpremultiply; compare where a is large
    big = want[..., 3] >= 128
    assert np.abs(got.astype(int) - want.astype(int))[big].max() <= 2


def test_convert_resize_handoff(cuda):
    """decoded Rgb8 buffer -> to_rgba8 on the device -> resize, one upload and one download"""
    rng = np.random.default_rng(150)
    rgb = rand_image(rng, 96, 128, 3, np.uint8)
    rgba = o.convert(rgb.reshape(-1, 3), 3, np.uint8, 4, o.CAST).reshape(96, 128, 4)
    want = o.resize(rgba, 64, 48, "lanczos3")
    got = imageops.convert_resize(rgb, 4, np.uint8, 64, 48, FilterType.Lanczos3)
    assert np.abs(got.astype(int) - want.astype(int)).max() <= 1
    from image_b200 import _lib as L
    strict = imageops.convert_resize(rgb, 4, np.uint8, 64, 48, FilterType.Lanczos3, flags=L.FP_STRICT)
    assert same(strict, want)
    # u16 -> f32 with a channel drop
    la16 = rand_image(rng, 50, 70, 2, np.uint16)
    l32 = o.convert(la16.reshape(-1, 2), 2, np.float32, 1, o.CAST).reshape(50, 70, 1)
    got = imageops.convert_resize(la16, 1, np.float32, 35, 25, FilterType.Triangle, flags=L.FP_STRICT)
    assert same(got, o.resize(l32, 35, 25, "triangle"))
