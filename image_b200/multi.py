"""Multi-GPU sharding of the hot path: one process per GPU, `torch.distributed` for the plumbing.

The path has no reduction across shards, so no collective is used:

* batches are sharded by image index (`shard_range`) and need no communication;
* one giant image is split into destination row bands (`band_plan`).  A band's vertical taps
  reach a few source rows into its neighbours' bands (the halo: `ceil(support * max(1, ratio)) + 1`
  rows per side, exactly the rows `b200img_resize_band_src_rows` reports), so before the kernel
  each rank fetches those rows from its neighbours with point-to-point sends (`exchange_halos`;
  NCCL send/recv over NVLink on GPUs, gloo on CPU for the tests).  The horizontal pass needs whole
  rows, so bands are never split in x.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import List

from . import _lib as L
from .imageops import FilterType, _call


def shard_range(n_items: int, world: int, rank: int) -> range:
    """Contiguous, balanced split of n_items over `world` ranks (first n % world ranks get one more)."""
    base, extra = divmod(n_items, world)
    start = rank * base + min(rank, extra)
    return range(start, start + base + (1 if rank < extra else 0))


@dataclass(frozen=True)
class Band:
    rank: int
    dst_row0: int   # destination rows [dst_row0, dst_row0 + dst_rows) computed by this rank
    dst_rows: int
    own_row0: int   # source rows this rank holds before the exchange (balanced split of h)
    own_rows: int
    src_row0: int   # source rows its destination band taps (own rows +- halo, clipped to the image)
    src_rows: int


def band_plan(h: int, nh: int, filter: FilterType, world: int) -> List[Band]:
    """Row-band decomposition of an h -> nh resize over `world` ranks."""
    lib = L.lib()
    bands = []
    for r in range(world):
        d = shard_range(nh, world, r)
        o = shard_range(h, world, r)
        s0, sn = C.c_uint32(), C.c_uint32()
        if len(d):
            _call(lib.b200img_resize_band_src_rows, h, nh, int(filter), d.start, len(d), C.byref(s0), C.byref(sn))
        bands.append(Band(r, d.start, len(d), o.start, len(o), s0.value, sn.value))
    return bands


def halo_transfers(bands: List[Band]):
    """[(src_rank, dst_rank, row0, rows)]: source rows dst_rank needs that src_rank owns."""
    out = []
    for b in bands:
        if not b.dst_rows:
            continue
        need0, need1 = b.src_row0, b.src_row0 + b.src_rows
        for o in bands:
            if o.rank == b.rank:
                continue
            lo, hi = max(need0, o.own_row0), min(need1, o.own_row0 + o.own_rows)
            if lo < hi:
                out.append((o.rank, b.rank, lo, hi - lo))
    return out


def exchange_halos(own_rows, bands: List[Band], rank: int, group=None):
    """own_rows: tensor (own_rows, w, c) holding this rank's source rows.  Returns the tensor of
    source rows [src_row0, src_row0 + src_rows) its destination band taps (own rows + halos)."""
    import torch
    import torch.distributed as dist
    me = bands[rank]
    out = torch.empty((me.src_rows,) + tuple(own_rows.shape[1:]), dtype=own_rows.dtype, device=own_rows.device)
    lo, hi = max(me.src_row0, me.own_row0), min(me.src_row0 + me.src_rows, me.own_row0 + me.own_rows)
    if lo < hi:
        out[lo - me.src_row0:hi - me.src_row0] = own_rows[lo - me.own_row0:hi - me.own_row0]
    ops, recvs = [], []
    for src, dst, r0, n in halo_transfers(bands):
        if src == rank:
            ops.append(dist.P2POp(dist.isend, own_rows[r0 - me.own_row0:r0 - me.own_row0 + n].contiguous(), dst, group))
        elif dst == rank:
            buf = torch.empty((n,) + tuple(own_rows.shape[1:]), dtype=own_rows.dtype, device=own_rows.device)
            recvs.append((r0, n, buf))
            ops.append(dist.P2POp(dist.irecv, buf, src, group))
    if ops:
        for req in dist.batch_isend_irecv(ops):
            req.wait()
    for r0, n, buf in recvs:
        out[r0 - me.src_row0:r0 - me.src_row0 + n] = buf
    return out


def resize_band(band_src, band: Band, w: int, h: int, nw: int, nh: int, filter: FilterType, *, acc_init: float = 1.0,
                flags: int = L.FP_FAST):
    """Runs b200img_resize_band_dev on this rank's band (CUDA tensor (src_rows, w, c)); returns (dst_rows, nw, c)."""
    import torch
    c = band_src.shape[2]
    dt = {torch.uint8: L.U8, torch.uint16: L.U16, torch.float32: L.F32}[band_src.dtype]
    out = torch.empty((band.dst_rows, nw, c), dtype=band_src.dtype, device=band_src.device)
    if band.dst_rows:
        with torch.cuda.device(band_src.device):
            _call(L.lib().b200img_resize_band_dev, band_src.data_ptr(), band.src_row0, band.src_rows, w, h, c, dt,
                  out.data_ptr(), band.dst_row0, band.dst_rows, nw, nh, int(filter), float(acc_init), flags,
                  torch.cuda.current_stream(band_src.device).cuda_stream)
    return out


# ---- the in-library split: one host thread drives every device (csrc/multi.cu) -------------------
def plan_bands(op: str, h: int, n_bands: int, *, nh: int = 0, filter: FilterType = FilterType.Nearest, params=None,
               sigma: float = 0.0):
    """(ctypes array of n_bands `_lib.Band`) row fields for an even split: op = 'resize' | 'gaussian' | 'fast_blur'."""
    lib = L.lib()
    arr = (L.Band * n_bands)()
    if op == "resize":
        _call(lib.b200img_plan_resize_bands, h, nh, int(filter), n_bands, arr)
    elif op == "gaussian":
        _call(lib.b200img_plan_gaussian_blur_bands, h, C.byref(params.raw), n_bands, arr)
    elif op == "fast_blur":
        _call(lib.b200img_plan_fast_blur_bands, h, float(sigma), n_bands, arr)
    else:
        raise ValueError(op)
    return arr


def run_bands(op: str, image, devices, *, nw: int = 0, nh: int = 0, filter: FilterType = FilterType.Nearest,
              acc_init: float = 1.0, flags: int = L.FP_FAST, params=None, path: int = L.BLUR_GENERIC, sigma: float = 0.0,
              ghost_fill: int = 0xAB):
    """Runs `op` on a host image (numpy (h, w, c)) split over `devices` (CUDA ordinals, one band each; an
    ordinal may repeat) through b200img_{resize,gaussian_blur,fast_blur}_bands.  Every device receives its
    OWN rows only; its ghost rows start as `ghost_fill` bytes and are filled by the library.  Returns the
    stitched result as a numpy array.  (A test / demo harness: a real caller keeps the bands resident.)"""
    import numpy as np
    import torch
    lib = L.lib()
    h, w, c = image.shape
    n = len(devices)
    oh = nh if op == "resize" else h
    ow = nw if op == "resize" else w
    bands = plan_bands(op, h, n, nh=nh, filter=filter, params=params, sigma=sigma)
    dt = {np.dtype(np.uint8): L.U8, np.dtype(np.uint16): L.U16, np.dtype(np.float32): L.F32}[image.dtype]
    keep = []
    for k, d in enumerate(devices):
        b = bands[k]
        dev = torch.device("cuda", d)
        src = torch.full((max(b.src_rows, 1) * w * c * image.dtype.itemsize,), ghost_fill, dtype=torch.uint8, device=dev)
        if b.own_rows:
            own = np.ascontiguousarray(image[b.own_row0:b.own_row0 + b.own_rows]).view(np.uint8).reshape(-1)
            off = (b.own_row0 - b.src_row0) * w * c * image.dtype.itemsize
            src[off:off + own.size] = torch.from_numpy(own).to(dev)
        dst = torch.zeros((max(b.dst_rows, 1) * ow * c * image.dtype.itemsize,), dtype=torch.uint8, device=dev)
        b.device, b.d_src, b.d_dst = d, src.data_ptr(), dst.data_ptr()
        keep.append((src, dst))
    for d in set(devices):
        torch.cuda.synchronize(d)
    if op == "resize":
        _call(lib.b200img_resize_bands, bands, n, w, h, c, dt, nw, nh, int(filter), float(acc_init), flags, None)
    elif op == "gaussian":
        _call(lib.b200img_gaussian_blur_bands, bands, n, w, h, c, dt, C.byref(params.raw), path, flags, None)
    else:
        _call(lib.b200img_fast_blur_bands, bands, n, w, h, c, dt, float(sigma), None)
    out = np.empty((oh, ow, c), image.dtype)
    for k in range(n):
        b = bands[k]
        if b.dst_rows:
            raw = keep[k][1].cpu().numpy()[:b.dst_rows * ow * c * image.dtype.itemsize]
            out[b.dst_row0:b.dst_row0 + b.dst_rows] = raw.view(image.dtype).reshape(b.dst_rows, ow, c)
    return out
