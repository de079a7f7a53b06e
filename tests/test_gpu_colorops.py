"""GPU parity for the elementwise colour operators: bit-exact against the oracle."""
import numpy as np
import pytest

import oracle as o
from conftest import rand_image, same
from image_b200 import imageops

pytestmark = pytest.mark.gpu
DT = [np.uint8, np.uint16, np.float32]


def both(fn_gpu, fn_cpu, img, cuda):
    """device-pointer and host-pointer entry points against the oracle"""
    import torch
    want = fn_cpu(img)
    got = fn_gpu(torch.from_numpy(img.copy()).to(cuda)).cpu().numpy()
    assert same(got, want), np.flatnonzero(got.ravel() != want.ravel())[:5]
    assert same(fn_gpu(img.copy()), want)


def test_reference_kats(cuda):
    """colorops.rs:657-700: test_grayscale, test_invert, test_brighten, test_brighten_place"""
    img = np.array([0, 1, 2, 10, 11, 12], np.uint8).reshape(2, 3, 1)
    assert imageops.grayscale(img).ravel().tolist() == [0, 1, 2, 10, 11, 12]
    assert imageops.brighten(img, 10).ravel().tolist() == [10, 11, 12, 20, 21, 22]
    b = img.copy()
    imageops.brighten_in_place(b, 10)
    assert b.ravel().tolist() == [10, 11, 12, 20, 21, 22]
    inv = img.copy()
    imageops.invert(inv)
    assert inv.ravel().tolist() == [255, 254, 253, 245, 244, 243]


@pytest.mark.parametrize("dtype", DT)
@pytest.mark.parametrize("c", [1, 2, 3, 4])
def test_colorops(cuda, dtype, c):
    rng = np.random.default_rng(100 + c)
    for (h, w) in ((37, 53), (64, 64), (1, 7)):
        img = rand_image(rng, h, w, c, dtype)
        both(lambda a: imageops.invert(a), o.invert, img, cuda)
        for v in (10, -40, 300, -70000, 2**31 - 1):
            both(lambda a: imageops.brighten(a, v), lambda a: o.brighten(a, v), img, cuda)
        for ct in (0.0, 20.0, -35.5, 250.0, -100.0):
            both(lambda a: imageops.contrast(a, ct), lambda a: o.contrast(a, ct), img, cuda)
        for deg in (0, 90, -15, 180, 345, 360):
            both(lambda a: imageops.huerotate(a, deg), lambda a: o.huerotate(a, deg), img, cuda)
        both(lambda a: imageops.grayscale(a), lambda a: o.grayscale(a), img, cuda)
        both(lambda a: imageops.grayscale_alpha(a), lambda a: o.grayscale(a, alpha=True), img, cuda)


def test_in_place_forms_and_unaligned_views(cuda):
    import torch
    rng = np.random.default_rng(110)
    img = rand_image(rng, 33, 47, 4, np.uint8)
    for fn, ref in ((lambda a: imageops.contrast_in_place(a, 30.0), lambda a: o.contrast(a, 30.0)),
                    (lambda a: imageops.huerotate_in_place(a, 77), lambda a: o.huerotate(a, 77)),
                    (lambda a: imageops.brighten_in_place(a, -20), lambda a: o.brighten(a, -20))):
        dev = torch.from_numpy(img.copy()).to(cuda)
        fn(dev)
        assert same(dev.cpu().numpy(), ref(img))
    # a device image that does not start on a 16-byte boundary takes the per-pixel path
    flat = torch.zeros(33 * 47 * 3 + 1, dtype=torch.uint8, device=cuda)
    rgb = rand_image(rng, 33, 47, 3, np.uint8)
    view = flat[1:].view(33, 47, 3)
    view.copy_(torch.from_numpy(rgb))
    assert same(imageops.huerotate(view, 120).cpu().numpy(), o.huerotate(rgb, 120))


def test_full_size_invert_involution(cuda):
    import torch
    g = torch.Generator(device="cpu").manual_seed(3)
    img = torch.randint(0, 256, (2160, 3840, 4), dtype=torch.uint8, generator=g)
    dev = img.to(cuda)
    imageops.invert(dev)
    once = dev.cpu()
    assert torch.equal(once[..., :3], 255 - img[..., :3]) and torch.equal(once[..., 3], img[..., 3])
    imageops.invert(dev)
    assert torch.equal(dev.cpu(), img)
