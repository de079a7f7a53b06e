"""b200img_fast_blur_noalloc_dev: the box blur for callers that must not allocate.

Some users of this library cannot let it call cudaMalloc: a serving process that has given
the whole GPU to its own memory pool, or a latency-critical loop that cannot absorb
cudaMalloc's device-wide synchronisation.  For them the library carries one extra entry point
whose transient image lives in a fixed 32 MiB array inside the library instead of an
allocated pool.  This test calls it through the C ABI and checks that it returns exactly
what the ordinary (allocating) entry point returns -- twice in a row, because the array is
shared between calls and keeps the previous call's contents -- and that it refuses images
that do not fit the array.

The file is named so that pytest collects it last: a run that stops at the first failure has
then already run every other test.
"""
import numpy as np
import pytest

from image_b200 import _lib as L

pytestmark = pytest.mark.gpu

NOALLOC_LIMIT = 32 << 20  # bytes; stated in include/b200img.h


def _both(lib, torch, img, sigma):
    """(no-alloc result, ordinary result) for one device image, on torch's current stream."""
    h, w, c = img.shape
    dt = {torch.uint8: L.U8, torch.uint16: L.U16, torch.float32: L.F32}[img.dtype]
    stream = torch.cuda.current_stream().cuda_stream
    got = torch.empty_like(img)
    want = torch.empty_like(img)
    L.check(lib.b200img_fast_blur_noalloc_dev(img.data_ptr(), w, h, c, dt, got.data_ptr(), sigma, stream))
    L.check(lib.b200img_fast_blur_dev(img.data_ptr(), w, h, c, dt, want.data_ptr(), sigma, stream))
    torch.cuda.synchronize()
    return got, want


def test_noalloc_matches_allocating_entry_point_twice_in_a_row(cuda):
    import torch
    lib = L.lib()
    g = torch.Generator(device=cuda).manual_seed(11)
    # two different images and radii back to back: the second call finds the first call's
    # leftovers in the shared array and must not be affected by them
    first = torch.randint(0, 256, (301, 517, 3), dtype=torch.uint8, device=cuda, generator=g)
    second = torch.randint(0, 256, (123, 211, 4), dtype=torch.uint8, device=cuda, generator=g)
    for img, sigma in ((first, 10.0), (second, 2.5), (first, 10.0)):
        got, want = _both(lib, torch, img, sigma)
        assert torch.equal(got, want)  # u8 box blur is integer arithmetic: bit-identical
    # the other sample types go through the same array
    f32 = torch.rand((97, 130, 4), device=cuda, generator=g)
    got, want = _both(lib, torch, f32, 3.0)
    assert torch.equal(got, want)
    u16 = torch.randint(0, 65536, (64, 80, 2), dtype=torch.int32, device=cuda, generator=g).to(torch.uint16)
    got, want = _both(lib, torch, u16, 4.0)
    assert torch.equal(got.view(torch.int16), want.view(torch.int16))


def test_noalloc_largest_image_and_size_limit(cuda):
    import torch
    lib = L.lib()
    stream = torch.cuda.current_stream().cuda_stream
    # 3840 x 2160 Rgba8 = 33,177,600 bytes: the largest common frame that fits
    big = torch.full((2160, 3840, 4), 143, dtype=torch.uint8, device=cuda)
    assert big.numel() <= NOALLOC_LIMIT
    out = torch.empty_like(big)
    L.check(lib.b200img_fast_blur_noalloc_dev(big.data_ptr(), 3840, 2160, 4, L.U8, out.data_ptr(), 10.0, stream))
    torch.cuda.synchronize()
    assert int(out.min()) == 143 and int(out.max()) == 143  # a constant image stays constant
    # one row more does not fit: refused with a message that names the limit, nothing is launched
    too_big = torch.zeros((2185, 3840, 4), dtype=torch.uint8, device=cuda)
    assert too_big.numel() > NOALLOC_LIMIT
    rc = lib.b200img_fast_blur_noalloc_dev(too_big.data_ptr(), 3840, 2185, 4, L.U8, too_big.data_ptr(), 10.0, stream)
    assert rc == L.ERR_UNSUPPORTED
    assert str(NOALLOC_LIMIT) in L.last_error()
    # argument errors are reported like the ordinary entry point's
    assert lib.b200img_fast_blur_noalloc_dev(big.data_ptr(), 8, 8, 5, L.U8, out.data_ptr(), 1.0, stream) == L.ERR_BAD_ARG
    assert lib.b200img_fast_blur_noalloc_dev(big.data_ptr(), 0, 0, 3, L.U8, out.data_ptr(), 1.0, stream) == 0


def test_noalloc_agrees_with_the_oracle(cuda):
    """Not only GPU against GPU: the reference's own arithmetic (CPU restatement) is the judge."""
    import torch
    import oracle
    lib = L.lib()
    rng = np.random.default_rng(23)
    stream = torch.cuda.current_stream().cuda_stream
    for (h, w, c), sigma in (((157, 203, 3), 4.0), ((64, 90, 4), 10.0)):
        img = rng.integers(0, 256, (h, w, c), dtype=np.uint8)
        dev = torch.from_numpy(img).to(cuda)
        out = torch.empty_like(dev)
        L.check(lib.b200img_fast_blur_noalloc_dev(dev.data_ptr(), w, h, c, L.U8, out.data_ptr(), sigma, stream))
        torch.cuda.synchronize()
        assert np.array_equal(out.cpu().numpy(), oracle.fast_blur(img, sigma))


def test_noalloc_really_allocates_nothing(cuda):
    """The property the entry point exists for.  A caller whose pool owns the whole GPU sees device
    memory in use; so after a first call (which only looks up the array's address) the free-memory
    figure of the driver must not move across further calls."""
    import torch
    lib = L.lib()
    # torch.cuda.mem_get_info is cudaMemGetInfo: the driver's figure, which counts every allocation of the
    # process whichever copy of the runtime made it (the library links its own, statically)
    stream = torch.cuda.current_stream().cuda_stream
    img = torch.randint(0, 256, (1201, 1603, 4), dtype=torch.uint8, device=cuda)
    out = torch.empty_like(img)
    L.check(lib.b200img_fast_blur_noalloc_dev(img.data_ptr(), 1603, 1201, 4, L.U8, out.data_ptr(), 6.0, stream))
    torch.cuda.synchronize()
    free_before, _ = torch.cuda.mem_get_info(cuda)
    for sigma in (6.0, 2.0, 15.0):
        L.check(lib.b200img_fast_blur_noalloc_dev(img.data_ptr(), 1603, 1201, 4, L.U8, out.data_ptr(), sigma, stream))
    torch.cuda.synchronize()
    free_after, _ = torch.cuda.mem_get_info(cuda)
    assert free_after == free_before, (free_before, free_after)
