"""Per-kernel shares of an ncu launch list (gpu__time_duration.sum csv)."""
import collections
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
hdr = rows[0]
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows[1:]:
    try:
        v = float(r[vi].replace(",", ""))
    except ValueError:
        continue
    scale = {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(r[ui], 1e-6)
    n = r[ki].split("(")[0].replace("void <unnamed>::", "")[:48]
    agg[n][0] += 1
    agg[n][1] += v * scale
tot = sum(v[1] for v in agg.values())
print(f"{'kernel':48s} {'launches':>8s} {'total ms':>10s} {'share':>7s}")
for n, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{n:48s} {c:8d} {t:10.3f} {100 * t / tot:6.1f}%")
