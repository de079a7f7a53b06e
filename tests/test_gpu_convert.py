"""GPU parity: conversions and swizzle through the C ABI, bit-exact against the oracle."""
import itertools

import numpy as np
import pytest

import oracle as o
from conftest import lcg_array, same
from image_b200 import DynamicImage, ImageBuffer, imageops
from image_b200 import _lib as L

pytestmark = pytest.mark.gpu
DT = [np.uint8, np.uint16, np.float32]


def _special(dtype, n):
    dtype = np.dtype(dtype)
    if dtype == np.float32:
        rng = np.random.default_rng(9)
        v = rng.random(n, dtype=np.float32) * 1.5 - 0.25
        k = np.arange(0, 256, dtype=np.float32)
        edge = np.concatenate([[np.nan, np.inf, -np.inf, -0.0, 0.0, 1.0, -1.0, 2.0, 1e-30, 0.5],
                               (k + 0.5) / 255.0, (k + 0.5) / 65535.0, np.nextafter((k + 0.5) / 255.0, 0).astype(np.float32)])
        v[:len(edge)] = edge[:n].astype(np.float32)
        return v
    return lcg_array(0x1234ABCD, n, dtype)


@pytest.mark.parametrize("kernel", ["tile", "direct", "auto"])
@pytest.mark.parametrize("mode", [L.CAST, L.CONVERT])
@pytest.mark.parametrize("sd,dd", list(itertools.product(DT, DT)))
def test_convert_all_layouts(cuda, sd, dd, mode, kernel, monkeypatch):
    """Every layout pair through both general conversion kernels (B200IMG_CONVERT_KERNEL forces one) and through the library's own choice."""
    import torch
    if kernel == "auto":   # the library's own choice, which includes the 16-pixel chunk kernel for Rgb8 sources
        monkeypatch.delenv("B200IMG_CONVERT_KERNEL", raising=False)
    else:
        monkeypatch.setenv("B200IMG_CONVERT_KERNEL", kernel)
    npx = 5000 + 37  # several tiles plus a ragged tail
    for sc, dc in itertools.product([1, 2, 3, 4], [1, 2, 3, 4]):
        src = _special(sd, npx * sc).reshape(npx, sc)
        want = o.convert(src, sc, dd, dc, mode)
        got = imageops.convert_pixels(torch.from_numpy(src).to(cuda), dc, dd, mode=mode).cpu().numpy()
        assert same(got, want), (sd, sc, dd, dc, mode)


def test_convert_exhaustive_integer_inputs(cuda):
    """Every u8 and u16 value through every per-channel rule (incl. the FMA-corrected division)."""
    import torch
    for sd, n in ((np.uint8, 256), (np.uint16, 65536)):
        src = np.arange(n, dtype=sd).reshape(-1, 1)
        for dd in DT:
            for mode in (L.CAST, L.CONVERT):
                for dc in (1, 3):
                    want = o.convert(src, 1, dd, dc, mode)
                    got = imageops.convert_pixels(torch.from_numpy(src).to(cuda), dc, dd, mode=mode).cpu().numpy()
                    assert same(got, want), (sd, dd, mode, dc)


def test_convert_host_pointer_path_and_coefs(cuda):
    src = _special(np.uint16, 3000 * 3).reshape(3000, 3)
    p3 = [0.22897343, 0.69173974, 0.07928687]
    got = imageops.convert_pixels(src, 2, np.uint8, coefs=p3)
    assert np.array_equal(got, o.convert(src, 3, np.uint8, 2, coefs=p3))


def test_convert_unaligned_and_tiny(cuda):
    import torch
    for npx in (1, 2, 1023, 1024, 1025, 2048):
        src = _special(np.uint8, npx * 3 + 1)
        dev = torch.from_numpy(src).to(cuda)[1:].reshape(npx, 3)  # base pointer off by one byte
        assert not dev.data_ptr() % 16 == 0
        lib = L.lib()
        out = torch.empty((npx, 4), dtype=torch.float32, device=cuda)
        L.check(lib.b200img_convert_dev(dev.data_ptr(), L.U8, 3, out.data_ptr(), L.F32, 4, npx, L.CAST, None, None))
        assert np.array_equal(out.cpu().numpy(), o.convert(src[1:].reshape(npx, 3), 3, np.float32, 4))


def test_k3_shapes_bit_exact(cuda):
    """BASELINE config 3 at a reduced size (full 8192x8192 runs in bench.py with the same check sampled)."""
    import torch
    n = 1024 * 1024
    a = lcg_array(0x1234ABCD, n * 4, np.uint16).reshape(n, 4)
    got = DynamicImage(ImageBuffer(torch.from_numpy(a).to(cuda).reshape(1024, 1024, 4))).to_rgb8().data.cpu().numpy()
    assert np.array_equal(got.reshape(n, 3), o.convert(a, 4, np.uint8, 3))
    b = lcg_array(0x1234ABCD, n * 3, np.uint8).reshape(n, 3)
    got = DynamicImage(ImageBuffer(torch.from_numpy(b).to(cuda).reshape(1024, 1024, 3))).to_rgba32f().data.cpu().numpy()
    assert np.array_equal(got.reshape(n, 4), o.convert(b, 3, np.float32, 4))


@pytest.mark.parametrize("dtype", DT)
@pytest.mark.parametrize("ch", [3, 4])
def test_swizzle(cuda, dtype, ch):
    import torch
    for n in (0, 1, ch - 1, ch, 4 * 1024 * ch, 4 * 1024 * ch + ch - 1, 4 * 1024 * ch + 5 * ch + 1, 100003, 1000003):
        a = _special(dtype, n) if n else np.zeros(0, dtype)
        dev = torch.from_numpy(a.copy()).to(cuda)
        imageops.swizzle_02(dev, ch)
        assert same(dev.cpu().numpy(), o.swizzle_02(a, ch)), (n, ch)
        host = a.copy()
        imageops.swizzle_02(host, ch)
        assert host.tobytes() == o.swizzle_02(a, ch).tobytes()


def test_swizzle_involution_large(cuda):
    """Size-independent property at the K3 size: swizzling twice is the identity; once swaps 0 and 2."""
    import torch
    n = 8192 * 8192 * 4
    g = torch.Generator(device=cuda).manual_seed(1)
    buf = torch.randint(0, 256, (n,), dtype=torch.uint8, device=cuda, generator=g)
    ref = buf.clone()
    imageops.swizzle_02(buf, 4)
    v, r = buf.view(-1, 4), ref.view(-1, 4)
    assert torch.equal(v[:, 0], r[:, 2]) and torch.equal(v[:, 2], r[:, 0]) and torch.equal(v[:, 1], r[:, 1]) and torch.equal(v[:, 3], r[:, 3])
    imageops.swizzle_02(buf, 4)
    assert torch.equal(buf, ref)


def test_from_raw_bgra_roundtrip(cuda):
    raw = lcg_array(5, 7 * 5 * 4, np.uint8)
    img = ImageBuffer.from_raw_bgra(7, 5, raw.copy())
    assert np.array_equal(img.data.reshape(-1), o.swizzle_02(raw, 4))
    assert np.array_equal(img.into_raw_bgra(), raw)
    assert ImageBuffer.from_raw_bgr(7, 5, raw[:10].copy()) is None


def test_convert_and_swizzle_split_in_ranges(cuda):
    """SURVEY 8(e), conversions / swizzle: elementwise, so an image may be split into any contiguous pixel ranges (one per GPU);
    the pieces, converted separately (on every visible device in turn), are bit-identical to the whole."""
    import torch
    lib = L.lib()
    rng = np.random.default_rng(5)
    npx = 1 << 20
    a = rng.integers(0, 65536, (npx, 4), dtype=np.uint16)
    want = o.convert(a, 4, np.uint8, 3)
    ndev = torch.cuda.device_count()
    cuts = [0, 1, 333_333, 700_001, npx]
    got = np.empty((npx, 3), np.uint8)
    sw = np.empty((npx, 4), np.uint16)
    for i in range(len(cuts) - 1):
        dev = torch.device("cuda", i % ndev)
        lo, hi = cuts[i], cuts[i + 1]
        with torch.cuda.device(dev):
            st = torch.cuda.current_stream().cuda_stream
            src = torch.from_numpy(a[lo:hi].copy()).to(dev)
            dst = torch.empty((hi - lo, 3), dtype=torch.uint8, device=dev)
            L.check(lib.b200img_convert_dev(src.data_ptr(), L.U16, 4, dst.data_ptr(), L.U8, 3, hi - lo, L.CAST, None, st))
            L.check(lib.b200img_swizzle_02_dev(src.data_ptr(), L.U16, 4, (hi - lo) * 4, st))
            got[lo:hi] = dst.cpu().numpy()
            sw[lo:hi] = src.cpu().numpy()
    assert np.array_equal(got, want)
    assert np.array_equal(sw, a[:, [2, 1, 0, 3]])
