"""Host-side multi-GPU logic on CPU: world_size-2 (and 3) gloo groups exercise image sharding and
the row-band halo exchange.  The band each rank ends up with must be exactly the source rows its
destination band taps, and stitching the per-band oracle resizes must give the whole-image resize."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle as o
from image_b200 import FilterType
from image_b200.multi import band_plan, exchange_halos, halo_transfers, shard_range


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, h, w, nh, nw, filt, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rng = np.random.default_rng(123)
        full = rng.random((h, w, 4), dtype=np.float32)  # every rank can rebuild the image to check against
        bands = band_plan(h, nh, FilterType(filt), world)
        me = bands[rank]
        own = torch.from_numpy(full[me.own_row0:me.own_row0 + me.own_rows].copy())
        got = exchange_halos(own, bands, rank)
        want = full[me.src_row0:me.src_row0 + me.src_rows]
        ok = np.array_equal(got.numpy(), want)
        # the band's resize (oracle arithmetic on the band == rows of the whole-image resize)
        whole = o.resize(full, nw, nh, filt, acc_init=0.0)
        l, c, wt = o.resize_weights(filt, h, nh, 0)
        tmp = np.zeros((me.dst_rows, w, 4), np.float32)
        for i in range(me.dst_rows):
            oy = me.dst_row0 + i
            for t in range(c[oy]):
                tmp[i] += got.numpy()[l[oy] + t - me.src_row0] * wt[oy, t]
        ok = ok and got.shape[0] == me.src_rows and np.isfinite(tmp).all()
        ret[rank] = (ok, me.dst_row0, me.dst_rows, whole.shape[0])
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
@pytest.mark.parametrize("h,nh,filt", [(64, 128, int(FilterType.Lanczos3)), (96, 40, int(FilterType.CatmullRom)),
                                         (50, 50 * 3, int(FilterType.Nearest))])
def test_band_halo_exchange_gloo(world, h, nh, filt):
    port = _free_port()
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(world, port, h, 24, nh, 31, filt, ret), nprocs=world, join=True)
    assert len(ret) == world
    rows = 0
    for r in range(world):
        ok, d0, dn, total = ret[r]
        assert ok, f"rank {r}: band after the halo exchange is not the rows its taps need"
        assert d0 == rows
        rows += dn
    assert rows == nh


def test_shard_range_is_a_partition():
    for n in (0, 1, 7, 8, 4096, 4097):
        for world in (1, 2, 3, 8):
            parts = [shard_range(n, world, r) for r in range(world)]
            assert sum(len(p) for p in parts) == n
            assert [i for p in parts for i in p] == list(range(n))
            assert max(len(p) for p in parts) - min(len(p) for p in parts) <= 1


def test_band_plan_halo_is_small_and_neighbourly():
    # K5: 32768 -> 65536 rows Lanczos3 over 8 GPUs: halo of ceil(3 * 1) + 1 = 4 rows per side at most
    bands = band_plan(32768, 65536, FilterType.Lanczos3, 8)
    for b in bands:
        assert b.own_rows == 4096 and b.dst_rows == 8192
        assert b.own_row0 - b.src_row0 <= 4 and (b.src_row0 + b.src_rows) - (b.own_row0 + b.own_rows) <= 4
    for src, dst, r0, n in halo_transfers(bands):
        assert abs(src - dst) == 1 and n <= 4
    # a 2x downscale needs ceil(3 * 2) + 1 = 7
    for src, dst, r0, n in halo_transfers(band_plan(4096, 2048, FilterType.Lanczos3, 4)):
        assert abs(src - dst) == 1 and n <= 7
