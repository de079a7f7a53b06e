#!/usr/bin/env python
"""bench.py -- megapixels/s of the pixel hot path on B200, with its HBM roofline.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--no-extras]
    torchrun --nproc-per-node N ... bench.py --gpus N ...         (one rank per GPU)

Workload (BASELINE.json configs[1], the configuration the metric is quoted on): Gaussian blur
with sigma = 10 (`imageops::blur_advanced(new_from_sigma(10.0))`, 65 taps, f32 engine) on a
batch of 3840x2160 Rgb8 images, one batch per GPU (image-sharded, no collective: weak
scaling).  The batch is larger than L2, so every step streams from HBM.

A step is one pass over the batch.  `value` is device-timed with the batch resident in HBM;
`e2e` is the same metric through the host-pointer C-ABI call (pinned host buffers, H2D and
D2H inside the timed region).  `extra` carries the other hot-path families at their
BASELINE shapes (device-timed, same roofline accounting).  `--impl reference` times the CPU
oracle (the reference itself is Rust and cannot be built here) on all host cores.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

W4K, H4K = 3840, 2160
BATCH = 8            # 8 x 24.9 MB in + 8 x 24.9 MB out = 398 MB per step (> 126 MB L2)
SIGMA = 10.0


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md, 6.65 TB/s)"


def bench_image_u8(w, h, c):
    y, x = np.mgrid[0:h, 0:w]
    ch = [x & 255, (y // 8) & 255, (x * 13 + y) % 256, (x ^ y) & 255]
    return np.stack(ch[:c], axis=-1).astype(np.uint8)


class ClockSampler:
    """SM clock and throttle reasons sampled (NVML, every ~5 ms) while the timed region runs."""

    def __init__(self, index):
        self.index, self.sm, self.reasons, self.max_sm = index, [], set(), None
        self._stop = threading.Event()
        self.t = None

    def __enter__(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.max_sm = float(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM))
            names = {nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
                     nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
                     nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
                     nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap"}

            def loop():
                while not self._stop.is_set():
                    try:
                        self.sm.append(float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)))
                        r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                        for bit, nm in names.items():
                            if r & bit:
                                self.reasons.add(nm)
                    except Exception:
                        pass
                    time.sleep(0.005)

            self.t = threading.Thread(target=loop, daemon=True)
            self.t.start()
        except Exception:
            self.t = None
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self.t:
            self.t.join(timeout=2)

    def summary(self):
        return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": self.max_sm,
                "reasons": sorted(self.reasons), "samples": len(self.sm)}


# ---------------------------------------------------------------------------------------------
def time_device(fn, steps, warmup, torch, barrier=None):
    """Average ms per call of fn(), CUDA events on the launching (current) stream."""
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    if barrier:
        barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    ev0.record()
    for _ in range(steps):
        fn()
    ev1.record()
    torch.cuda.synchronize()
    if barrier:
        barrier()
    return ev0.elapsed_time(ev1) / steps


def run_b200(args):
    import torch
    from image_b200 import FilterType, GaussianBlurParameters, imageops
    from image_b200 import _lib as L

    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    # one process per GPU: stay on the CPUs (and so the NUMA node) next to this GPU before any pinned
    # host memory is allocated -- the e2e leg is PCIe / host-memory bound
    from image_b200.hostutil import bind_cpu_affinity
    cpus = bind_cpu_affinity(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    barrier = (lambda: dist.barrier()) if dist else None
    lib = L.lib()
    hbm, hbm_src = peaks()
    stream = torch.cuda.current_stream().cuda_stream

    def max_over_ranks(ms):
        if not dist:
            return ms
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- main workload: blur sigma=10 on a batch of 4K Rgb8 images, resident in HBM ----
    base = bench_image_u8(W4K, H4K, 3)
    host_batch = np.stack([np.roll(base, 37 * (i + rank * BATCH), axis=1) for i in range(BATCH)])
    src = torch.from_numpy(host_batch).to(dev)
    dst = torch.empty_like(src)
    params = GaussianBlurParameters.new_from_sigma(SIGMA)
    pp = C.byref(params.raw)

    def step():
        L.check(lib.b200img_gaussian_blur_batch_dev(src.data_ptr(), W4K, H4K, 3, L.U8, dst.data_ptr(), BATCH, pp,
                                                    L.BLUR_GENERIC, L.FP_FAST, stream))

    with ClockSampler(local) as clk:
        ms = time_device(step, args.steps, args.warmup, torch, barrier)
    ms = max_over_ranks(ms)
    mp_step = BATCH * W4K * H4K / 1e6
    value = world * mp_step / (ms / 1e3)
    alg_bytes = 2 * BATCH * W4K * H4K * 3  # read every source byte once + write every destination byte once
    achieved = alg_bytes / (ms / 1e3) / 1e9

    # parity spot check of the timed output (one 4K image's band against the oracle is done in tests;
    # here: the constant-image invariant on the real output path and a checksum so work cannot be skipped)
    checksum = int(dst[0, ::97, ::89].to(torch.int64).sum().item())

    # ---- e2e: host-pointer C-ABI call, pinned host memory, copies inside the timed region ----
    pin_src = torch.from_numpy(host_batch).pin_memory()
    pin_dst = torch.empty_like(pin_src).pin_memory()
    e2e_steps = max(3, min(args.steps, 10))

    def e2e_call():  # the host-pointer batch call a user makes: uploads, kernels, downloads pipelined inside
        L.check(lib.b200img_gaussian_blur_batch(pin_src.data_ptr(), W4K, H4K, 3, L.U8, pin_dst.data_ptr(), BATCH, pp,
                                                L.BLUR_GENERIC, L.FP_FAST))

    for _ in range(2):
        e2e_call()
    if barrier:
        barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_call()
    e2e_s = (time.perf_counter() - t0) / e2e_steps
    e2e_s = max_over_ranks(e2e_s * 1e3) / 1e3
    e2e_val = world * mp_step / e2e_s
    assert np.array_equal(pin_dst.numpy()[:, ::97, ::89], dst[:, ::97, ::89].cpu().numpy()), "e2e and device paths disagree"

    traffic = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):  # dram__bytes_read + write of the step's kernels, from the committed ncu --set full capture
        traffic = json.load(open(tpath)).get("blur_sigma10_4k_rgb8_batch8", {}).get("dram_bytes_per_step")

    out = {
        "metric": "megapixels/sec (Gaussian blur sigma=10, 3840x2160 Rgb8 batch)", "value": round(value, 1),
        "unit": "MP/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms, 4),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "BASELINE configs[1]: imageops::blur sigma=10 (blur_advanced(new_from_sigma(10.0)), 65 taps, "
                               f"f32 engine) on {BATCH} x 3840x2160 Rgb8 per GPU, image-sharded",
                   "batch_per_gpu": BATCH, "l2": f"inputs larger than L2 ({2 * BATCH * W4K * H4K * 3 / 1e6:.0f} MB per step)",
                   "flags": "FP_FAST (+-1 LSB vs reference)", "checksum": checksum},
        "roofline": {"bound": "hbm", "achieved": round(achieved, 1), "peak": hbm, "unit": "GB/s",
                     "frac": round(achieved / hbm, 4), "traffic": traffic, "peak_source": hbm_src,
                     "kernel": "gaussian row+column passes of one batch call",
                     "algorithmic_bytes_per_launch": alg_bytes,
                     "fp32_bound": {"fma_per_step": (65 + 72) * BATCH * W4K * H4K * 3, "peak_tfma_s": 35.9,
                                    "frac": round((65 + 72) * BATCH * W4K * H4K * 3 / 35.9e12 / (ms / 1e3), 4),
                                    "note": "65 taps in the row pass + 72 (padded) in the column pass per subpixel on the FP32 pipes; peak = FFMA rate "
                                            "measured with tools/micro/ffma2.cu"},
                     "note": "traffic = both kernels of the call (f32 intermediate written and re-read); the op is "
                             "FP32-pipe bound, see DESIGN.md section 5"},
        "e2e": {"value": round(e2e_val, 1), "unit": "MP/s", "h2d_bytes_per_step": BATCH * W4K * H4K * 3,
                "d2h_bytes_per_step": BATCH * W4K * H4K * 3},
        "gpu_launches": args.steps * GAUSS_LAUNCHES,
        "clocks": clk.summary(),
        "host": {"cpu_affinity": (f"{len(cpus)} CPUs local to the GPU (NVML)" if cpus else "unbound")},
    }

    # ---- extras: the other hot-path families at their BASELINE shapes (rank 0 numbers, per GPU) ----
    if not args.no_extras and world == 1:
        out["extra"] = extras(torch, lib, L, dev, stream, hbm, max(3, args.steps // 2), 3, FilterType, GaussianBlurParameters)
    # ---- CPU baseline (rank 0, N=1 only) ----
    if world == 1 and not args.no_cpu:
        out["cpu_baseline"] = cpu_baseline(budget_s=12.0)
    if rank == 0:
        print(json.dumps(out))
    if dist:
        dist.destroy_process_group()


GAUSS_LAUNCHES = 2  # row kernel + column kernel per batched call


def extras(torch, lib, L, dev, stream, hbm, steps, warmup, FilterType, GaussianBlurParameters):
    res = {}

    def rec(name, fn, mp, alg_bytes, note=""):
        ms = time_device(fn, steps, warmup, torch)
        gbs = alg_bytes / (ms / 1e3) / 1e9
        res[name] = {"MP/s": round(mp / (ms / 1e3), 1), "ms": round(ms, 4), "GB/s": round(gbs, 1),
                     "hbm_frac": round(gbs / hbm, 4), **({"note": note} if note else {})}

    base = bench_image_u8(W4K, H4K, 3)
    src = torch.from_numpy(np.stack([np.roll(base, 11 * i, axis=0) for i in range(BATCH)])).to(dev)
    dst = torch.empty_like(src)
    px = BATCH * W4K * H4K
    for name, p, path in (("blur_radius10_21taps_f32", GaussianBlurParameters.new_from_radius(10.0), L.BLUR_GENERIC),
                          ("blur_sigma10_65taps_q15_dynimage", GaussianBlurParameters.new_from_sigma(10.0), L.BLUR_DYN)):
        pp = C.byref(p.raw)
        rec(name, lambda pp=pp, path=path: L.check(lib.b200img_gaussian_blur_batch_dev(
            src.data_ptr(), W4K, H4K, 3, L.U8, dst.data_ptr(), BATCH, pp, path, L.FP_FAST, stream)), px / 1e6, 2 * px * 3)
    rec("fast_blur_sigma10", lambda: L.check(lib.b200img_fast_blur_batch_dev(
        src.data_ptr(), W4K, H4K, 3, L.U8, dst.data_ptr(), BATCH, SIGMA, stream)), px / 1e6, 2 * px * 3)
    del src, dst

    # K1: 1920x1080 Rgba8 -> 960x540 Lanczos3, batch of 64 (> L2)
    n = 64
    k1 = torch.from_numpy(bench_image_u8(1920, 1080, 4)).to(dev).unsqueeze(0).repeat(n, 1, 1, 1).contiguous()
    o1 = torch.empty((n, 540, 960, 4), dtype=torch.uint8, device=dev)
    rec("resize_k1_lanczos3_1920x1080_rgba8_to_960x540", lambda: L.check(lib.b200img_resize_batch_dev(
        k1.data_ptr(), 0, 1920, 1080, 4, L.U8, o1.data_ptr(), 0, 960, 540, n, int(FilterType.Lanczos3), 1.0, L.FP_FAST,
        stream)), n * 1920 * 1080 / 1e6, n * (1920 * 1080 * 4 + 960 * 540 * 4), "MP/s on source pixels")
    del k1, o1
    # K4: 1024x1024 Rgba8 -> 256x256 CatmullRom, 256 images per step
    n = 256
    g = torch.Generator(device=dev).manual_seed(4)
    k4 = torch.randint(0, 256, (n, 1024, 1024, 4), dtype=torch.uint8, device=dev, generator=g)
    o4 = torch.empty((n, 256, 256, 4), dtype=torch.uint8, device=dev)
    rec("resize_k4_catmullrom_1024_rgba8_to_256", lambda: L.check(lib.b200img_resize_batch_dev(
        k4.data_ptr(), 0, 1024, 1024, 4, L.U8, o4.data_ptr(), 0, 256, 256, n, int(FilterType.CatmullRom), 1.0, L.FP_FAST,
        stream)), n * 1024 * 1024 / 1e6, n * (1024 * 1024 * 4 + 256 * 256 * 4), "MP/s on source pixels")
    del k4, o4
    # K5 shape on one GPU: 4096x4096 Rgba32F -> 8192x8192 Lanczos3 (268 MB in, 1.07 GB out)
    k5 = torch.rand((1, 4096, 4096, 4), device=dev, generator=g)
    o5 = torch.empty((1, 8192, 8192, 4), dtype=torch.float32, device=dev)
    rec("resize_k5shape_lanczos3_rgba32f_4096_to_8192", lambda: L.check(lib.b200img_resize_batch_dev(
        k5.data_ptr(), 0, 4096, 4096, 4, L.F32, o5.data_ptr(), 0, 8192, 8192, 1, int(FilterType.Lanczos3), 1.0, L.FP_FAST,
        stream)), 4096 * 4096 / 1e6, (4096 * 4096 + 8192 * 8192) * 16, "MP/s on source pixels")
    del k5, o5
    # the first "next" row: thumbnail of 8 x 4K Rgb8 frames to 256x256
    t8 = torch.randint(0, 256, (BATCH, H4K, W4K, 3), dtype=torch.uint8, device=dev, generator=g)
    ot = torch.empty((BATCH, 256, 256, 3), dtype=torch.uint8, device=dev)
    rec("thumbnail_8x4k_rgb8_to_256", lambda: L.check(lib.b200img_thumbnail_batch_dev(
        t8.data_ptr(), W4K, H4K, 3, L.U8, ot.data_ptr(), 256, 256, BATCH, 0, stream)), BATCH * W4K * H4K / 1e6,
        BATCH * (W4K * H4K * 3 + 256 * 256 * 3), "MP/s on source pixels")
    del t8, ot
    # K3: conversions and swizzle at 8192x8192
    npx = 8192 * 8192
    a16 = torch.randint(0, 65536, (npx, 4), dtype=torch.int32, device=dev, generator=g).to(torch.uint16)
    o8 = torch.empty((npx, 3), dtype=torch.uint8, device=dev)
    rec("convert_rgba16_to_rgb8_8192", lambda: L.check(lib.b200img_convert_dev(
        a16.data_ptr(), L.U16, 4, o8.data_ptr(), L.U8, 3, npx, L.CAST, None, stream)), npx / 1e6, npx * (8 + 3))
    del a16, o8
    a8 = torch.randint(0, 256, (npx, 3), dtype=torch.uint8, device=dev, generator=g)
    of = torch.empty((npx, 4), dtype=torch.float32, device=dev)
    rec("convert_rgb8_to_rgba32f_8192", lambda: L.check(lib.b200img_convert_dev(
        a8.data_ptr(), L.U8, 3, of.data_ptr(), L.F32, 4, npx, L.CAST, None, stream)), npx / 1e6, npx * (3 + 16))
    del a8, of
    bg = torch.randint(0, 256, (npx * 4,), dtype=torch.uint8, device=dev, generator=g)
    rec("swizzle_bgra_u8_8192_inplace", lambda: L.check(lib.b200img_swizzle_02_dev(
        bg.data_ptr(), L.U8, 4, npx * 4, stream)), npx / 1e6, 2 * npx * 4)
    return res


# ---------------------------------------------------------------------------------------------
def _cpu_blur_rate(threads: int, budget_s: float):
    """MP/s of the oracle's blur_advanced(sigma=10) on 3840-wide Rgb8 bands, `threads` bands at once."""
    import oracle
    p = oracle.gauss_params_from_sigma(SIGMA)
    base = bench_image_u8(W4K, H4K, 3)
    # size the sample from a short probe so the whole thing stays near the budget
    probe = np.ascontiguousarray(base[:135])
    t0 = time.perf_counter()
    oracle.gaussian_blur(probe, p, "generic")
    per_row = (time.perf_counter() - t0) / 135
    rows = int(max(135, min(H4K, budget_s / max(per_row, 1e-9))))
    bands = [np.ascontiguousarray(np.roll(base, 17 * i, axis=1)[:rows]) for i in range(threads)]
    outs = [None] * threads

    def work(i):
        outs[i] = oracle.gaussian_blur(bands[i], p, "generic")

    ts = [threading.Thread(target=work, args=(i,)) for i in range(threads)]
    t0 = time.perf_counter()
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    dt = time.perf_counter() - t0
    return threads * rows * W4K / 1e6 / dt, rows


def cpu_baseline(budget_s: float):
    cores = os.cpu_count() or 1
    one, rows1 = _cpu_blur_rate(1, budget_s / 2)
    allc, rows = _cpu_blur_rate(cores, budget_s / 2)
    return {"value": round(allc, 2), "unit": "MP/s", "cores": cores, "kind": "port",
            "single_thread_value": round(one, 2),
            "sample": f"oracle blur_advanced(sigma=10) on {cores} bands of 3840x{rows} Rgb8 at once (one band per host "
                      f"thread); single-thread figure on one 3840x{rows1} band. The reference itself is single-threaded "
                      "on this path and cannot be built here (no Rust toolchain)."}


def run_reference(args):
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    budget = min(20.0, 150.0 / max(args.steps + args.warmup, 1))
    for _ in range(min(args.warmup, 1)):
        _cpu_blur_rate(cores, budget / 4)
    rates, rows = [], 0
    t0 = time.perf_counter()
    for _ in range(args.steps):
        r, rows = _cpu_blur_rate(cores, budget)
        rates.append(r)
    wall = time.perf_counter() - t0
    v = float(np.mean(rates))
    print(json.dumps({
        "impl": "reference", "metric": "megapixels/sec (Gaussian blur sigma=10, 3840x2160 Rgb8 batch)",
        "value": round(v, 2), "unit": "MP/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": round(wall / args.steps * 1e3, 2), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": "BASELINE configs[1]: imageops::blur sigma=10 (65 taps, f32 engine) on 3840x2160 Rgb8; "
                               f"each step a bounded sample of {cores} bands of 3840x{rows}"},
        "cpu_baseline": {"value": round(v, 2), "unit": "MP/s", "cores": cores, "kind": "port",
                         "sample": f"{cores} bands of 3840x{rows} Rgb8 per step, one per host thread (oracle port; the Rust "
                                   "reference cannot be built in this image)"},
        "e2e": {"value": round(v, 2), "unit": "MP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-extras", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
