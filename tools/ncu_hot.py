"""Per-instruction hot spots from an .ncu-rep source page.  usage: ncu_hot.py REP KERNEL_REGEX [top]"""
import csv, io, subprocess, sys, collections
rep, pat = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{pat}"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hi = [i for i, r in enumerate(rows) if "Source" in r and "# Samples" in r]
h = rows[hi[0]]; ix = {c: i for i, c in enumerate(h)}
end = hi[1] - 1 if len(hi) > 1 else len(rows)
data = [r for r in rows[hi[0] + 1:end] if len(r) == len(h)]
S = lambda r: int(r[ix["# Samples"]] or 0); I = lambda r: int(r[ix["Instructions Executed"]] or 0)
ts, ti = sum(map(S, data)), sum(map(I, data))
print(f"kernel {pat}: sass={len(data)} samples={ts} warp-inst={ti}")
for r in sorted(data, key=lambda r: -S(r))[:top]:
    print(f"  {r[ix['Address']][-5:]} {100*S(r)/ts:5.1f}%  inst={I(r):10d}  {r[ix['Source']][:80]}")
a0 = int(data[0][ix["Address"]], 16)
b = collections.OrderedDict()
for r in data:
    k = (int(r[ix["Address"]], 16) - a0) // 0x400
    b.setdefault(k, [0, 0]); b[k][0] += S(r); b[k][1] += I(r)
print("  region   samples%  inst%")
for k, v in b.items():
    if v[0] or v[1]:
        print(f"  {k*0x400:#7x}  {100*v[0]/ts:6.1f}  {100*v[1]/ti:6.1f}")
