"""GPU parity for imageops::thumbnail against the oracle, through the C ABI."""
import numpy as np
import pytest

import oracle as o
from conftest import rand_image, same
from image_b200 import imageops
from image_b200 import _lib as L

pytestmark = pytest.mark.gpu
DT = [np.uint8, np.uint16, np.float32]


def check(img, nw, nh, cuda, *, paths=(0, L.THUMBNAIL_PER_PIXEL)):
    import torch
    want = o.thumbnail(img, nw, nh)
    dev = torch.from_numpy(img).to(cuda)
    for flags in paths:
        got = imageops.thumbnail(dev, nw, nh, flags=flags).cpu().numpy()
        if img.dtype == np.float32:
            # f64 sums in another order, rounded to f32 once: 1e-6 relative (bar: 1e-5)
            np.testing.assert_allclose(got, want, rtol=1e-6, atol=1e-7, err_msg=f"flags={flags}")
        else:
            assert same(got, want), (img.shape, nw, nh, flags, np.abs(got.astype(int) - want.astype(int)).max())


@pytest.mark.parametrize("dtype", DT)
@pytest.mark.parametrize("c", [1, 2, 3, 4])
def test_downscale(cuda, dtype, c):
    rng = np.random.default_rng(80 + c)
    img = rand_image(rng, 96, 128, c, dtype)      # rows are word-aligned: column-sum kernel
    for nw, nh in ((64, 48), (13, 7), (128, 96), (100, 90), (1, 1), (5, 96), (128, 3)):
        check(img, nw, nh, cuda)


@pytest.mark.parametrize("dtype", DT)
def test_unaligned_rows_and_odd_sizes(cuda, dtype):
    rng = np.random.default_rng(90)
    for (h, w, c) in ((57, 61, 3), (33, 35, 1), (41, 23, 2)):
        img = rand_image(rng, h, w, c, dtype)
        for nw, nh in ((20, 19), (7, 5), (w, h), (w - 1, h - 1)):
            check(img, nw, nh, cuda)


@pytest.mark.parametrize("dtype", DT)
def test_upscale_and_mixed(cuda, dtype):
    """ratio < 1 on an axis: the fraction_horizontal / _vertical / _both forms (sample.rs:630-670)"""
    rng = np.random.default_rng(91)
    img = rand_image(rng, 24, 32, 3, dtype)
    for nw, nh in ((64, 48), (45, 37), (33, 25), (100, 10), (10, 100), (32, 70), (200, 24), (31, 25)):
        check(img, nw, nh, cuda)
    one = rand_image(rng, 1, 1, 4, dtype)
    check(one, 5, 3, cuda)


def test_wide_image_many_segments(cuda):
    rng = np.random.default_rng(92)
    img = rand_image(rng, 40, 5000, 3, np.uint8)   # 15000 subpixels per row: several x segments
    for nw, nh in ((2500, 20), (333, 13), (4999, 40), (3, 2)):
        check(img, nw, nh, cuda)
    img16 = rand_image(rng, 16, 6000, 4, np.uint16)
    check(img16, 1000, 4, cuda)
    f = rand_image(rng, 16, 3000, 3, np.float32)
    check(f, 700, 5, cuda)


def test_windows_above_65536_pixels(cuda):
    """sample.rs:695-716: 256x256 sub-block averages, divided by (w/256)*(h/256) -- literally"""
    rng = np.random.default_rng(93)
    img = rng.integers(0, 100, (700, 600, 1), dtype=np.uint8)
    check(img, 1, 1, cuda)               # 3x3 sub-blocks over a count of 2*2
    check(img, 2, 1, cuda)               # 300x700: 2x3 sub-blocks over 1*2
    img16 = rng.integers(0, 20000, (520, 530, 2), dtype=np.uint16)
    check(img16, 1, 1, cuda)
    f = rng.random((520, 530, 1), dtype=np.float32)
    check(f, 1, 1, cuda)


def test_reference_kat_thumbnail_huge_block(cuda):
    """sample.rs:1922-1930: must not overflow / panic"""
    import torch
    huge_u8 = torch.full((4113, 4096, 1), 255, dtype=torch.uint8, device=cuda)
    assert int(imageops.thumbnail(huge_u8, 1, 1).cpu().numpy()[0, 0, 0]) == 255
    huge_u16 = torch.full((1024, 1024, 1), 65535, dtype=torch.uint16, device=cuda)
    assert int(imageops.thumbnail(huge_u16, 1, 1).cpu().numpy()[0, 0, 0]) == 65535


def test_reference_panic_is_an_error(cuda):
    thin = np.zeros((1000, 100, 1), np.uint8)
    with pytest.raises(o.ReferencePanic):
        o.thumbnail(thin, 1, 1)
    with pytest.raises(imageops.PanicError):
        imageops.thumbnail(thin, 1, 1)
    f = np.full((1000, 100, 1), 0.5, np.float32)   # f32: x / 0.0 = inf -> f32::MAX, no panic
    assert same(imageops.thumbnail(f, 1, 1), o.thumbnail(f, 1, 1))


def test_empty_and_batch(cuda):
    import torch
    assert not imageops.thumbnail(np.zeros((0, 5, 3), np.uint8), 4, 4).any()
    assert imageops.thumbnail(np.zeros((6, 5, 3), np.uint8), 0, 4).shape == (4, 0, 3)
    rng = np.random.default_rng(94)
    batch = np.stack([rand_image(rng, 64, 80, 3, np.uint8) for _ in range(3)])
    got = imageops.thumbnail(torch.from_numpy(batch).to(cuda), 20, 16).cpu().numpy()
    for i in range(3):
        assert same(got[i], o.thumbnail(batch[i], 20, 16))
    # host-pointer entry point
    assert same(imageops.thumbnail(batch[0], 33, 21), o.thumbnail(batch[0], 33, 21))


def test_full_size_exact_ratio_property(cuda):
    """4K frame to 384x216 (ratio exactly 10): every output is the rounded mean of its 10x10 block"""
    import torch
    g = torch.Generator(device="cpu").manual_seed(5)
    img = torch.randint(0, 256, (2, 2160, 3840, 3), dtype=torch.uint8, generator=g)
    got = imageops.thumbnail(img.to(cuda), 384, 216).cpu().numpy()
    s = img.numpy().reshape(2, 216, 10, 384, 10, 3).astype(np.uint32).sum(axis=(2, 4))
    assert np.array_equal(got, ((s + 50) // 100).astype(np.uint8))
