"""Warp-sample and instruction shares per 0x200-byte code region of a kernel, with the top stall reasons.
usage: ncu_regions.py REP KERNEL_REGEX"""
import collections
import csv
import io
import subprocess
import sys

rep, pat = sys.argv[1], sys.argv[2]
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{pat}"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hi = [i for i, r in enumerate(rows) if "Source" in r and "# Samples" in r]
h = rows[hi[0]]
ix = {c: i for i, c in enumerate(h)}
end = hi[1] - 1 if len(hi) > 1 else len(rows)
data = [r for r in rows[hi[0] + 1:end] if len(r) == len(h)]
reasons = [c for c in h if c.startswith("stall_") and "Not Issued" not in c]
base = int(data[0][ix["Address"]], 16)
tot = sum(int(r[ix["# Samples"]] or 0) for r in data)
b = collections.OrderedDict()
for r in data:
    a = (int(r[ix["Address"]], 16) - base) // 0x200
    n = int(r[ix["# Samples"]] or 0)
    ie = int(r[ix["Instructions Executed"]] or 0)
    e = b.setdefault(a, [0, 0, collections.Counter()])
    e[0] += n
    e[1] += ie
    for c in reasons:
        v = r[ix[c]]
        if v:
            e[2][c[6:]] += int(v)
ti = sum(v[1] for v in b.values())
print(f"samples {tot}, warp instructions {ti}")
for k, v in b.items():
    if v[0] > tot * 0.004:
        print(f"{k * 0x200:#7x} samples {100 * v[0] / tot:5.1f}% inst {100 * v[1] / ti:5.1f}%  {dict(v[2].most_common(3))}")
