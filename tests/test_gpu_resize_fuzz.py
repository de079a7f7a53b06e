"""Seeded sweep of odd shapes through the resampler's kernel selection (streaming, group, tile, unfused):
STRICT results are bit-exact against the oracle, FAST within 1 LSB / 1e-5."""
import numpy as np
import pytest

import oracle as o
from conftest import rand_image, same
from image_b200 import FilterType, imageops
from image_b200 import _lib as L

pytestmark = pytest.mark.gpu
FILTERS = [(FilterType.Nearest, "nearest"), (FilterType.Triangle, "triangle"), (FilterType.CatmullRom, "catmullrom"),
           (FilterType.Gaussian, "gaussian"), (FilterType.Lanczos3, "lanczos3")]


@pytest.mark.parametrize("seed", range(6))
def test_random_shapes(cuda, seed):
    import torch
    rng = np.random.default_rng(1000 + seed)
    dims = [1, 2, 3, 4, 5, 7, 8, 15, 16, 17, 31, 33, 64, 65, 100, 127, 130, 257]
    for _ in range(28):
        h, w, nh, nw = (int(rng.choice(dims)) for _ in range(4))
        c = int(rng.integers(1, 5))
        dtype = [np.uint8, np.uint16, np.float32][int(rng.integers(0, 3))]
        f, fname = FILTERS[int(rng.integers(0, 5))]
        acc = float(rng.choice([1.0, 0.0]))
        img = rand_image(rng, h, w, c, dtype)
        want = o.resize(img, nw, nh, fname, acc_init=acc)
        dev = torch.from_numpy(img).to(cuda)
        tag = (h, w, c, np.dtype(dtype).name, nw, nh, fname, acc)
        strict = imageops.resize(dev, nw, nh, f, acc_init=acc, flags=L.FP_STRICT).cpu().numpy()
        assert same(strict, want), tag
        fast = imageops.resize(dev, nw, nh, f, acc_init=acc).cpu().numpy()
        if dtype == np.float32:
            np.testing.assert_allclose(fast, want, rtol=1e-5, atol=1e-6, err_msg=str(tag))
        else:
            assert np.abs(fast.astype(np.int64) - want.astype(np.int64)).max() <= 1, tag


def test_group_kernel_bands_and_batches(cuda):
    """upscales through the band entry point (rows aligned to four and not) and a strided batch"""
    import torch
    rng = np.random.default_rng(77)
    img = rand_image(rng, 40, 52, 4, np.uint8)
    want = o.resize(img, 104, 80, "catmullrom")
    lib = L.lib()
    dev = torch.from_numpy(img).to(cuda)
    st = torch.cuda.current_stream().cuda_stream
    for (r0, rows) in ((0, 80), (16, 32), (18, 30), (44, 36), (3, 9)):
        s0, sn = L.C.c_uint32(), L.C.c_uint32()
        L.check(lib.b200img_resize_band_src_rows(40, 80, int(FilterType.CatmullRom), r0, rows, L.C.byref(s0), L.C.byref(sn)))
        band = dev[s0.value:s0.value + sn.value].contiguous()
        out = torch.empty((rows, 104, 4), dtype=torch.uint8, device=cuda)
        L.check(lib.b200img_resize_band_dev(band.data_ptr(), s0.value, sn.value, 52, 40, 4, L.U8, out.data_ptr(), r0, rows,
                                            104, 80, int(FilterType.CatmullRom), 1.0, L.FP_STRICT, st))
        assert same(out.cpu().numpy(), want[r0:r0 + rows]), (r0, rows)
    batch = np.stack([rand_image(rng, 30, 44, 3, np.uint16) for _ in range(3)])
    got = imageops.resize_batch(torch.from_numpy(batch).to(cuda), 90, 61, FilterType.Lanczos3, flags=L.FP_STRICT).cpu().numpy()
    for i in range(3):
        assert same(got[i], o.resize(batch[i], 90, 61, "lanczos3")), i
