"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: time per kernel name over one step."""
import csv, sys, collections, re
path, first, count = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
rows = []
with open(path) as f:
    lines = [l for l in f if not l.startswith("==")]
for r in csv.DictReader(lines):
    if r.get("Metric Name") == "gpu__time_duration.sum":
        v = float(r["Metric Value"].replace(",", ""))
        unit = r["Metric Unit"]
        ns = v * {"ns": 1, "us": 1e3, "usecond": 1e3, "ms": 1e6, "msecond": 1e6, "nsecond": 1, "second": 1e9}.get(unit, 1)
        rows.append((int(r["ID"]), r["Kernel Name"], ns))
rows = rows[first:first + count]
agg = collections.defaultdict(lambda: [0, 0.0])
for _, k, ns in rows:
    k = re.sub(r"\(.*", "", k)
    agg[k][0] += 1; agg[k][1] += ns
tot = sum(v[1] for v in agg.values())
print(f"launches {len(rows)} total {tot/1e6:.3f} ms")
for k, (n, ns) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{ns/1e6:9.3f} ms {100*ns/tot:5.1f}% n={n:4d}  {k[:110]}")
