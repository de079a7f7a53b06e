"""The host <-> device copy ceiling with EVERY rank copying at once (one process per GPU, pinned memory allocated
after binding to the GPU's local CPUs, both directions busy, 24.9 MB chunks = one 4K Rgb8 frame): what bounds the
end-to-end line at N GPUs.   torchrun --nproc-per-node N tools/pcie_bw_ranks.py"""
import os
import sys
import time

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from image_b200.hostutil import bind_cpu_affinity  # noqa: E402

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
bind_cpu_affinity(local)
if world > 1:
    dist.init_process_group("gloo")
dev = torch.device("cuda", local)
CH, NCH = 24883200, 16
h_in = torch.empty(CH * NCH, dtype=torch.uint8).pin_memory()
h_out = torch.empty(CH * NCH, dtype=torch.uint8).pin_memory()
d_a = torch.empty(CH * NCH, dtype=torch.uint8, device=dev)
d_b = torch.empty(CH * NCH, dtype=torch.uint8, device=dev)
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def run(h2d, d2h):
    best = 1e9
    for _ in range(3):
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        for i in range(NCH):
            sl = slice(i * CH, (i + 1) * CH)
            if h2d:
                with torch.cuda.stream(s1):
                    d_a[sl].copy_(h_in[sl], non_blocking=True)
            if d2h:
                with torch.cuda.stream(s2):
                    h_out[sl].copy_(d_b[sl], non_blocking=True)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        if world > 1:
            t = torch.tensor([dt], dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt = float(t.item())
        best = min(best, dt)
    return CH * NCH / best / 1e9


a, b, c = run(True, False), run(False, True), run(True, True)
if rank == 0:
    print(f"{world} ranks at once, per rank (time = slowest rank): H2D alone {a:.1f} GB/s, D2H alone {b:.1f} GB/s, both at once {c:.1f} GB/s each way; "
          f"aggregate both-at-once {c * world:.1f} GB/s each way")
if world > 1:
    dist.destroy_process_group()
