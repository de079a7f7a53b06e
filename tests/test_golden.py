"""Golden vectors (tests/golden/digests.json, written by tests/golden/make_golden.py from the oracle):
the oracle must keep reproducing them, and the CUDA path must produce the same bytes without the
oracle in the loop."""
import json
import os
import sys

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
from cases import cases, digest  # noqa: E402

with open(os.path.join(HERE, "golden", "digests.json")) as f:
    GOLDEN = json.load(f)


def test_golden_file_covers_every_case():
    names = [name for name, *_ in cases()]
    assert len(names) == len(set(names))
    assert set(names) == set(GOLDEN)


def test_oracle_reproduces_golden_vectors():
    bad = [name for name, img, oracle_fn, _ in cases() if digest(oracle_fn(img)) != GOLDEN[name]]
    assert not bad, bad[:10]


@pytest.mark.gpu
def test_cuda_reproduces_golden_vectors(cuda):
    import numpy as np
    bad, ran = [], 0
    for name, img, _, gpu_fn in cases():
        if gpu_fn is None:
            continue
        got = np.asarray(gpu_fn(img.copy()))
        ran += 1
        if digest(got) != GOLDEN[name]:
            bad.append(name)
    assert ran >= 150 and not bad, (ran, bad[:10])
