"""Aggregate an `ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,
sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active --csv` log over one training step:
per kernel name: launches, time, DRAM bytes, achieved DRAM GB/s, time-weighted tensor-pipe activity.
    python scripts/step_metrics.py <csv> <first launch> <count>"""
import collections
import csv
import re
import sys

path, first, count = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
lines = [l for l in open(path) if not l.startswith("==")]
per = collections.OrderedDict()
scale = {"ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1.0, "nsecond": 1e-9, "usecond": 1e-6, "msecond": 1e-3, "second": 1.0,
         "byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "%": 1.0}
for r in csv.DictReader(lines):
    k = int(r["ID"])
    d = per.setdefault(k, {"name": re.sub(r"\(.*", "", r["Kernel Name"])})
    v = float(r["Metric Value"].replace(",", "")) * scale.get(r["Metric Unit"], 1.0)
    d[r["Metric Name"]] = v
ids = sorted(per)[first:first + count]
agg = collections.defaultdict(lambda: [0, 0.0, 0.0, 0.0])
for k in ids:
    d = per[k]
    t = d.get("gpu__time_duration.sum", 0.0)
    a = agg[d["name"]]
    a[0] += 1
    a[1] += t
    a[2] += d.get("dram__bytes_read.sum", 0.0) + d.get("dram__bytes_write.sum", 0.0)
    a[3] += t * d.get("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", 0.0)
tot = sum(a[1] for a in agg.values())
print(f"launches {len(ids)}  kernel time {tot*1e3:.3f} ms  DRAM traffic {sum(a[2] for a in agg.values())/1e9:.2f} GB")
print(f"{'ms':>8} {'share':>6} {'n':>4} {'DRAM MB':>9} {'GB/s':>7} {'tensor%':>8}  kernel")
for name, (n, t, b, tp) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{t*1e3:8.3f} {100*t/tot:5.1f}% {n:4d} {b/1e6:9.1f} {b/t/1e9 if t else 0:7.0f} {tp/t if t else 0:8.1f}  {name[:90]}")
