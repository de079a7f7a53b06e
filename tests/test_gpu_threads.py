"""Host threads sharing the library: two threads on two streams drive the resampler (enough distinct
shapes to turn over the per-device weight-table cache), both blurs (whose transient images come from
the scratch pool) and the host-pointer entry points at once; every result must equal the one the
same call gives when it runs alone.  ctypes drops the GIL for the duration of a call, so the calls
really overlap."""
import threading

import numpy as np
import pytest

from image_b200 import FilterType, GaussianBlurParameters, imageops
from image_b200 import _lib as L

pytestmark = pytest.mark.gpu


def _jobs(rng, k):
    jobs = []
    for i in range(k):
        h, w = int(rng.integers(40, 90)), int(rng.integers(40, 120))
        nh, nw = int(rng.integers(20, 130)), int(rng.integers(20, 130))
        img = rng.integers(0, 256, (h, w, 3 + (i & 1)), dtype=np.uint8)
        jobs.append((img, nw, nh, FilterType(1 + i % 4), 1.0 + (i % 5)))
    return jobs


def _run(job, torch, dev, host):
    img, nw, nh, filt, sigma = job
    p = GaussianBlurParameters.new_from_sigma(sigma)
    if host:
        return (imageops.resize(img, nw, nh, filt, flags=L.FP_STRICT), imageops.blur_dyn(img, p), imageops.fast_blur(img, sigma))
    d = torch.from_numpy(img).to(dev)
    out = (imageops.resize(d, nw, nh, filt, flags=L.FP_STRICT), imageops.blur_dyn(d, p), imageops.fast_blur(d, sigma))
    return tuple(o.cpu().numpy() for o in out)


@pytest.mark.parametrize("host", [False, True], ids=["device-pointers", "host-pointers"])
def test_two_threads_two_streams(cuda, host):
    import torch
    rng = np.random.default_rng(5)
    jobs = _jobs(rng, 48)  # 96 (filter, length) keys per axis: more than the 32-entry cache holds
    want = [_run(j, torch, cuda, host) for j in jobs]
    L.check(L.lib().b200img_release_caches())
    got = [None] * len(jobs)
    errors = []

    def worker(tid):
        try:
            with torch.cuda.device(cuda), torch.cuda.stream(torch.cuda.Stream(cuda)):
                for i in range(tid, len(jobs), 2):
                    got[i] = _run(jobs[i], torch, cuda, host)
                torch.cuda.current_stream().synchronize()
        except Exception as e:  # noqa: BLE001
            errors.append(e)

    for _ in range(2):
        ts = [threading.Thread(target=worker, args=(t,)) for t in range(2)]
        for t in ts:
            t.start()
        for t in ts:
            t.join()
        assert not errors, errors
        for g, w in zip(got, want):
            for a, b in zip(g, w):
                assert np.array_equal(a, b)
