"""Two-GPU row-band resize: NCCL send/recv halo exchange over NVLink, then b200img_resize_band_dev
per rank; the stitched result must equal the whole-image oracle resize.  Skipped on 1-GPU boxes."""
import os
import socket

import numpy as np
import pytest

import oracle as o
from image_b200 import FilterType
from image_b200 import _lib as L

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, h, w, nh, nw, ret):
    import torch
    import torch.distributed as dist
    from image_b200.multi import band_plan, exchange_halos, resize_band
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        rng = np.random.default_rng(77)
        full = rng.random((h, w, 4), dtype=np.float32)
        bands = band_plan(h, nh, FilterType.Lanczos3, world)
        me = bands[rank]
        own = torch.from_numpy(full[me.own_row0:me.own_row0 + me.own_rows].copy()).cuda()
        band_src = exchange_halos(own, bands, rank)
        out = resize_band(band_src, me, w, h, nw, nh, FilterType.Lanczos3, acc_init=0.0, flags=L.FP_STRICT)
        torch.cuda.synchronize()
        ret[rank] = (me.dst_row0, out.cpu().numpy())
        dist.barrier()
    finally:
        dist.destroy_process_group()


def test_two_gpu_band_resize_matches_oracle(cuda):
    import torch
    import torch.multiprocessing as mp
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    h, w, nh, nw = 200, 96, 400, 192
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(2, _free_port(), h, w, nh, nw, ret), nprocs=2, join=True)
    rng = np.random.default_rng(77)
    full = rng.random((h, w, 4), dtype=np.float32)
    want = o.resize(full, nw, nh, "lanczos3", acc_init=0.0)
    got = np.concatenate([ret[r][1] for r in sorted(ret.keys(), key=lambda r: ret[r][0])], axis=0)
    assert got.tobytes() == want.tobytes()
