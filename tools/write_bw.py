"""What can HBM take when the traffic is mostly stores?  torch fill_ (stores only), copy_ (1:1) and a broadcast copy (1 read : 4 written)."""
import torch


def t(fn, n=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


h = w = 8192
src = torch.rand((h, w, 4), device="cuda")
dst = torch.empty((2 * h, 2 * w, 4), device="cuda")
big = torch.empty_like(dst)
ms = t(lambda: dst.fill_(1.0))
print(f"fill_ {dst.numel() * 4 / 1e9:.2f} GB: {ms:.3f} ms = {dst.numel() * 4 / ms / 1e6:.0f} GB/s")
ms = t(lambda: big.copy_(dst))
print(f"copy_ {2 * dst.numel() * 4 / 1e9:.2f} GB moved: {ms:.3f} ms = {2 * dst.numel() * 4 / ms / 1e6:.0f} GB/s")
d5 = dst.view(h, 2, w, 2, 4)
s5 = src.view(h, 1, w, 1, 4)
ms = t(lambda: d5.copy_(s5.expand(h, 2, w, 2, 4)))
alg = (src.numel() + dst.numel()) * 4
print(f"broadcast copy 1:4, {alg / 1e9:.2f} GB: {ms:.3f} ms = {alg / ms / 1e6:.0f} GB/s")
