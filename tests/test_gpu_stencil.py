"""GPU parity for the first "next" rows: filter3x3 and unsharpen against the oracle."""
import numpy as np
import pytest

import oracle as o
from conftest import rand_image, same
from image_b200 import imageops
from image_b200 import _lib as L

pytestmark = pytest.mark.gpu
DT = [np.uint8, np.uint16, np.float32]
KERNELS = {
    "identity": [0, 0, 0, 0, 1, 0, 0, 0, 0],
    "box": [1] * 9,
    "sharpen": [0, -1, 0, -1, 5, -1, 0, -1, 0],
    "zero_sum_edge": [-1, -1, -1, -1, 8, -1, -1, -1, -1],   # sum 0 -> inverse_sum = 1
    "frac": [0.0625, 0.125, 0.0625, 0.125, 0.25, 0.125, 0.0625, 0.125, 0.0625],
}


@pytest.mark.parametrize("dtype", DT)
@pytest.mark.parametrize("c", [1, 2, 3, 4])
def test_filter3x3(cuda, dtype, c):
    import torch
    rng = np.random.default_rng(60 + c)
    img = rand_image(rng, 37, 53, c, dtype)
    dev = torch.from_numpy(img).to(cuda)
    for name, k in KERNELS.items():
        want = o.filter3x3(img, k)
        assert same(imageops.filter3x3(dev, k).cpu().numpy(), want), name
    assert same(imageops.filter3x3(img, KERNELS["frac"]), o.filter3x3(img, KERNELS["frac"]))  # host-pointer path
    # the border keeps the zero fill (sample.rs:921: loops run over 1..h-1, 1..w-1)
    out = imageops.filter3x3(img, KERNELS["box"])
    assert not out[0].any() and not out[-1].any() and not out[:, 0].any() and not out[:, -1].any()


def test_filter3x3_tiny_images(cuda):
    for shape in ((1, 1, 3), (2, 9, 1), (9, 2, 4), (3, 3, 2)):
        img = np.full(shape, 200, np.uint8)
        assert same(imageops.filter3x3(img, KERNELS["box"]), o.filter3x3(img, KERNELS["box"])), shape


@pytest.mark.parametrize("dtype", DT)
def test_unsharpen(cuda, dtype):
    import torch
    rng = np.random.default_rng(70)
    for c in (1, 3, 4):
        img = rand_image(rng, 41, 57, c, dtype)
        dev = torch.from_numpy(img).to(cuda)
        for sigma, thr in ((1.5, 0), (3.0, 5), (0.1, 2)):
            want = o.unsharpen(img, sigma, thr)
            got = imageops.unsharpen(dev, sigma, thr, flags=L.FP_STRICT).cpu().numpy()
            assert same(got, want), (c, sigma, thr)
            if dtype != np.float32:
                # FAST blur is within 1 LSB, so the mask differs by at most 1 except where the +-1 moves
                # |diff| across the threshold (then by at most threshold + 2)
                fast = imageops.unsharpen(dev, sigma, thr).cpu().numpy().astype(np.int64)
                d = np.abs(fast - want.astype(np.int64))
                assert d.max() <= thr + 2 and (d > 1).mean() < 0.02
    with pytest.raises(imageops.PanicError):
        imageops.unsharpen(np.zeros((4, 4, 3), np.uint8), -1.0, 1)
