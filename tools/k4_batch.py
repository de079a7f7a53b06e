"""BASELINE config 4: 4096 images 1024x1024 Rgba8 -> 256x256 CatmullRom, device-resident, image-sharded
over the ranks (no communication).  torchrun --nproc-per-node N tools/k4_batch.py [--images 4096] [--steps 5]
Job time = max over ranks of the CUDA-event time for the rank's shard.  Prints one JSON line."""
import argparse, json, os, sys
import torch
import torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from image_b200 import FilterType
from image_b200 import _lib as L
from image_b200.multi import shard_range


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--images", type=int, default=4096)
    ap.add_argument("--steps", type=int, default=5)
    a = ap.parse_args()
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    shard = shard_range(a.images, world, rank)
    lo, n = shard.start, len(shard)
    g = torch.Generator(device=dev).manual_seed(1 + lo)
    src = torch.randint(0, 256, (n, 1024, 1024, 4), dtype=torch.uint8, device=dev, generator=g)
    dst = torch.empty((n, 256, 256, 4), dtype=torch.uint8, device=dev)
    lib, st = L.lib(), torch.cuda.current_stream().cuda_stream

    def step():
        L.check(lib.b200img_resize_batch_dev(src.data_ptr(), 0, 1024, 1024, 4, L.U8, dst.data_ptr(), 0, 256, 256, n,
                                             int(FilterType.CatmullRom), 1.0, L.FP_FAST, st))

    for _ in range(3):
        step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / a.steps
    if world > 1:
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    alg = a.images * (1024 * 1024 * 4 + 256 * 256 * 4)
    if rank == 0:
        print(json.dumps({"workload": f"K4 {a.images} x 1024x1024 Rgba8 -> 256x256 CatmullRom, image-sharded over {world} GPU(s)",
                          "n_gpus": world, "ms": round(ms, 3), "src_MP/s": round(a.images * 1.048576 / (ms / 1e3), 1),
                          "algorithmic_GB/s_total": round(alg / (ms / 1e3) / 1e9, 1),
                          "hbm_frac_per_gpu": round(alg / (ms / 1e3) / 1e9 / world / 6555.5, 4),
                          "checksum": int(dst[::max(n // 7, 1), ::17, ::13].to(torch.int64).sum().item())}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
