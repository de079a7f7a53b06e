"""Band plans (host arithmetic only, no GPU): every source row a band reads is owned by exactly one
band, owned rows lie inside the owner's buffer, destination bands tile the output."""
import ctypes as C

import numpy as np
import pytest

from image_b200 import FilterType, GaussianBlurParameters
from image_b200 import _lib as L
from image_b200.multi import plan_bands


def _check(bands, h, dst_h):
    owner = np.full(h, -1)
    nxt = 0
    for k, b in enumerate(bands):
        assert b.dst_row0 == nxt
        nxt += b.dst_rows
        assert b.src_row0 + b.src_rows <= h
        if b.own_rows:
            assert b.src_row0 <= b.own_row0 and b.own_row0 + b.own_rows <= b.src_row0 + b.src_rows
            assert (owner[b.own_row0:b.own_row0 + b.own_rows] == -1).all()
            owner[b.own_row0:b.own_row0 + b.own_rows] = k
    assert nxt == dst_h
    for b in bands:
        assert (owner[b.src_row0:b.src_row0 + b.src_rows] >= 0).all()


@pytest.mark.parametrize("n", [1, 2, 3, 8, 13])
def test_resize_plans(n):
    lib = L.lib()
    for filt in FilterType:
        for h, nh in ((32768, 65536), (1080, 540), (100, 7), (7, 100), (50, 50), (5, 3)):
            bands = plan_bands("resize", h, n, nh=nh, filter=filt)
            _check(bands, h, nh)
            for b in bands:
                if b.dst_rows:
                    s0, sn = C.c_uint32(), C.c_uint32()
                    L.check(lib.b200img_resize_band_src_rows(h, nh, int(filt), b.dst_row0, b.dst_rows, C.byref(s0), C.byref(sn)))
                    assert (s0.value, sn.value) == (b.src_row0, b.src_rows)
                    assert b.dst_row0 % 4 == 0


@pytest.mark.parametrize("n", [1, 2, 8, 11])
def test_blur_plans(n):
    for h in (2160, 100, 9):
        p = GaussianBlurParameters.new_from_sigma(10.0)
        bands = plan_bands("gaussian", h, n, params=p)
        _check(bands, h, h)
        for b in bands:
            if b.dst_rows:
                assert b.src_row0 == max(0, b.dst_row0 - 32) and b.src_row0 + b.src_rows == min(h, b.dst_row0 + b.dst_rows + 32)
        bands = plan_bands("fast_blur", h, n, sigma=10.0)   # radii 9 + 9 + 11
        _check(bands, h, h)
        for b in bands:
            if b.dst_rows:
                assert b.src_row0 == max(0, b.dst_row0 - 29) and b.src_row0 + b.src_rows == min(h, b.dst_row0 + b.dst_rows + 29)
