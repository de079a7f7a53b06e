"""Writes tests/golden/digests.json from the CPU oracle.  usage: python tests/golden/make_golden.py
(Re-run only when the oracle is deliberately changed; the diff of digests.json then shows what moved.)"""
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, HERE)
from cases import cases, digest  # noqa: E402

out = {}
for name, img, oracle_fn, _ in cases():
    assert name not in out, name
    out[name] = digest(oracle_fn(img))
with open(os.path.join(HERE, "digests.json"), "w") as f:
    json.dump(out, f, indent=0, sort_keys=True)
print(len(out), "cases")
