"""Sustained run of the headline kernel: the batched Gaussian blur back to back for a few seconds, with the
SM clock, power draw and throttle reasons sampled (NVML, every 20 ms) and the per-step time of every 100-step
window.  Shows whether the clock holds under this load.   python tools/sustained.py [seconds] [flags]"""
import ctypes as C
import os
import sys
import threading
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from image_b200 import GaussianBlurParameters  # noqa: E402
from image_b200 import _lib as L  # noqa: E402

secs = float(sys.argv[1]) if len(sys.argv) > 1 else 3.0
flags = int(sys.argv[2]) if len(sys.argv) > 2 else 0
lib = L.lib()
dev = torch.device("cuda:0")
st = torch.cuda.current_stream().cuda_stream
src = torch.randint(0, 256, (8, 2160, 3840, 3), dtype=torch.uint8, device=dev)
dst = torch.empty_like(src)
p = GaussianBlurParameters.new_from_sigma(10.0)


def step():
    L.check(lib.b200img_gaussian_blur_batch_dev(src.data_ptr(), 3840, 2160, 3, L.U8, dst.data_ptr(), 8, C.byref(p.raw), L.BLUR_GENERIC, flags, st))


import pynvml as nv  # noqa: E402
nv.nvmlInit()
h = nv.nvmlDeviceGetHandleByIndex(0)
samples, stop = [], threading.Event()


def sampler():
    while not stop.is_set():
        samples.append((time.perf_counter(), nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM), nv.nvmlDeviceGetPowerUsage(h) / 1000.0,
                        nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)))
        time.sleep(0.02)


for _ in range(5):
    step()
torch.cuda.synchronize()
t = threading.Thread(target=sampler, daemon=True)
t.start()
t0 = time.perf_counter()
windows = []
while time.perf_counter() - t0 < secs:
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(100):
        step()
    e1.record()
    torch.cuda.synchronize()
    windows.append((time.perf_counter() - t0, e0.elapsed_time(e1) / 100))
stop.set()
t.join()
print("flags", flags, "ms/step per 100-step window:", [round(w[1], 4) for w in windows])
clk = [s[1] for s in samples]
pw = [s[2] for s in samples]
reasons = 0
for s in samples:
    reasons |= s[3]
print(f"sm clock MHz: min {min(clk)} median {int(np.median(clk))} max {max(clk)}; power W: median {np.median(pw):.0f} max {max(pw):.0f}; throttle reason bits {reasons:#x} "
      f"(0x4 = sw_power_cap, 0x8 = hw_slowdown, 0x20 = sw_thermal, 0x40 = hw_thermal)")
print("clock trace (every 10th sample):", clk[::10])
