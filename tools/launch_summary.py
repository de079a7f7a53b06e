"""Summarises an ncu launch list (--metrics gpu__time_duration.sum --csv --log-file F).  usage: launch_summary.py F.csv"""
import collections, csv, re, sys
rows = list(csv.reader(l for l in open(sys.argv[1]) if not l.startswith("==")))
h = rows[0]
ik, iv = h.index("Kernel Name"), h.index("Metric Value")
agg = collections.OrderedDict()
for r in rows[1:]:
    if len(r) != len(h):
        continue
    name = re.sub(r"\(.*", "", r[ik])
    a = agg.setdefault(name, [0, 0.0])
    a[0] += 1
    a[1] += float(r[iv].replace(",", "")) / 1e3
tot = sum(v[1] for v in agg.values())
print(f"{'kernel':92s} {'launches':>8s} {'total_us':>10s} {'avg_us':>9s}  share")
for k, (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{k[-92:]:92s} {n:8d} {us:10.1f} {us / n:9.1f} {100 * us / tot:5.1f}%")
