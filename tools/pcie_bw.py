"""Pinned host <-> device copy bandwidth on this box: each direction alone, both at once, and by chunk size.
The end-to-end (host-resident) bench lines are bounded by these numbers."""
import torch, time
dev = torch.device("cuda:0")
N = 512 << 20
h_in = torch.empty(N, dtype=torch.uint8).pin_memory()
h_out = torch.empty(N, dtype=torch.uint8).pin_memory()
d_a = torch.empty(N, dtype=torch.uint8, device=dev)
d_b = torch.empty(N, dtype=torch.uint8, device=dev)
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def run(h2d, d2h, chunk, reps=3):
    best = 1e9
    for _ in range(reps):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for off in range(0, N, chunk):
            if h2d:
                with torch.cuda.stream(s1):
                    d_a[off:off + chunk].copy_(h_in[off:off + chunk], non_blocking=True)
            if d2h:
                with torch.cuda.stream(s2):
                    h_out[off:off + chunk].copy_(d_b[off:off + chunk], non_blocking=True)
        torch.cuda.synchronize()
        best = min(best, time.perf_counter() - t0)
    return N / best / 1e9


for chunk in (N, 64 << 20, 24883200, 8 << 20, 2 << 20):
    print(f"chunk {chunk / 2**20:7.1f} MiB: H2D alone {run(True, False, chunk):6.1f} GB/s   D2H alone {run(False, True, chunk):6.1f} GB/s   "
          f"both at once {run(True, True, chunk):6.1f} GB/s each way")
