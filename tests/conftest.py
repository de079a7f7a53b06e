import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session", autouse=True)
def _built():
    """Both libraries are built once per session (nvcc and gcc work without a GPU)."""
    from image_b200 import build as b
    b.build()
    import oracle
    oracle.build()


@pytest.fixture(scope="session")
def cuda():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch.device("cuda:0")


class Lcg:
    """The reference's own test RNG (src/imageops/fast_blur.rs:519-522)."""

    def __init__(self, seed):
        self.state = seed & 0xFFFFFFFFFFFFFFFF

    def next_u32(self):
        self.state = (self.state * 6364136223846793005 + 1) & 0xFFFFFFFFFFFFFFFF
        return self.state >> 32


def lcg_array(seed, n, dtype):
    """n values from the LCG (vectorised: x_k = a^k x_0 + c_k mod 2^64)."""
    # fall back to numpy's generator seeded from the LCG for large n (the values only need
    # to be deterministic and full-range; small cases use the literal sequence)
    if n <= 4096:
        g = Lcg(seed)
        v = np.array([g.next_u32() for _ in range(n)], dtype=np.uint64)
    else:
        v = np.random.default_rng(Lcg(seed).next_u32()).integers(0, 2 ** 32, n, dtype=np.uint64)
    dtype = np.dtype(dtype)
    if dtype == np.uint8:
        return (v & 0xFF).astype(np.uint8)
    if dtype == np.uint16:
        return (v & 0xFFFF).astype(np.uint16)
    return (v.astype(np.float64) / 2.0 ** 32).astype(np.float32)


def bench_image(w, h, c):
    """The reference's bench generator (benches/imageops.rs:11-14), alpha = (x ^ y) & 255."""
    y, x = np.mgrid[0:h, 0:w]
    chans = [x & 255, (y // 8) & 255, (x * 13 + y) % 256, (x ^ y) & 255]
    return np.stack(chans[:c], axis=-1).astype(np.uint8)


def rand_image(rng, h, w, c, dtype):
    dtype = np.dtype(dtype)
    if dtype == np.float32:
        return rng.random((h, w, c), dtype=np.float32)
    return rng.integers(0, np.iinfo(dtype).max + 1, (h, w, c), dtype=dtype)


def same(got, want):
    """Bit-identical, except that any NaN matches any NaN: an x86 host produces the 'default NaN'
    0xFFC00000 where the GPU produces 0x7FFFFFFF; which NaN comes out is not part of the
    reference's contract (and differs between the CPU targets it supports)."""
    got, want = np.asarray(got), np.asarray(want)
    if got.shape != want.shape or got.dtype != want.dtype:
        return False
    if got.dtype.kind != "f":
        return got.tobytes() == want.tobytes()
    nan = np.isnan(want)
    if not np.array_equal(np.isnan(got), nan):
        return False
    gi, wi = got.view(np.uint32), want.view(np.uint32)
    return bool(np.array_equal(gi[~nan], wi[~nan]))
