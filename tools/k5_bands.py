"""BASELINE config 5 (single giant Rgba32F image, Lanczos3 2x upscale, row-band split with a halo
exchange over NVLink), at a size given on the command line.  Launch with torchrun, one rank per GPU:

    torchrun --nproc-per-node N tools/k5_bands.py [--side 32768] [--steps 5]

Every rank generates its own source band on the device (uniform [0,1) from a counter hash, seed 7),
fetches the halo rows from its neighbours with NCCL send/recv, resizes its band, and the job time is
the max over ranks of the CUDA-event time (halo exchange included).  Prints one JSON line."""
import argparse, json, os, sys
import torch
import torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from image_b200 import FilterType
from image_b200 import _lib as L
from image_b200.multi import band_plan, exchange_halos, resize_band


def gen_rows(row0, rows, w, device):
    """uniform [0,1) from a counter hash of (row, x, c); identical on every rank for the same coordinates"""
    idx = (torch.arange(row0, row0 + rows, device=device, dtype=torch.int64)[:, None, None] * w
           + torch.arange(w, device=device, dtype=torch.int64)[None, :, None]) * 4 + torch.arange(4, device=device)[None, None, :]
    x = (idx * 0x9E3779B97F4A7C15 + 7) & 0x7FFFFFFFFFFFFFFF
    x = (x ^ (x >> 31)) * 0x7FB5D329728EA185 & 0x7FFFFFFFFFFFFFFF
    x = x ^ (x >> 27)
    return ((x & 0xFFFFFF).to(torch.float32) / 16777216.0).contiguous()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--side", type=int, default=32768)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--acc-init", type=float, default=0.0)
    a = ap.parse_args()
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    h = w = a.side
    nh, nw = 2 * h, 2 * w
    bands = band_plan(h, nh, FilterType.Lanczos3, world)
    me = bands[rank]
    own = gen_rows(me.own_row0, me.own_rows, w, dev)

    def step():
        src = exchange_halos(own, bands, rank) if world > 1 else own
        return resize_band(src, me, w, h, nw, nh, FilterType.Lanczos3, acc_init=a.acc_init)

    out = step()
    out = step()   # twice: the caching allocator then holds both output buffers the timed loop alternates between
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        out = step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / a.steps
    if world > 1:
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    # property check at full size: with acc_init = 0 and inputs in [0,1) the Lanczos3 overshoot is bounded
    ok = bool(torch.isfinite(out).all()) and float(out.min()) >= 0.0 and float(out.max()) <= 1.0
    alg = h * w * 16 + nh * nw * 16
    if rank == 0:
        print(json.dumps({"workload": f"K5 {h}x{w} Rgba32F -> {nh}x{nw} Lanczos3, {world} row bands, NCCL halo exchange",
                          "n_gpus": world, "ms": round(ms, 3), "src_MP/s": round(h * w / 1e6 / (ms / 1e3), 1),
                          "algorithmic_GB/s_total": round(alg / (ms / 1e3) / 1e9, 1),
                          "hbm_frac_per_gpu": round(alg / (ms / 1e3) / 1e9 / world / 6555.5, 4),
                          "halo_rows_per_edge": max([b.own_row0 - b.src_row0 for b in bands] + [0]), "clamped_output_ok": ok}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
