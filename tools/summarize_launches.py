"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: time per kernel name, share of the step."""
import csv
import collections
import re
import sys

path, per_step = sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 1
rows = []
with open(path, newline="") as fh:
    lines = [l for l in fh if not l.startswith("==")]
for r in csv.DictReader(lines):
    if r.get("Metric Name") == "gpu__time_duration.sum":
        v = float(r["Metric Value"].replace(",", ""))
        unit = r.get("Metric Unit", "ns")
        v = v / 1e3 if unit in ("ns", "nsecond") else (v if unit in ("us", "usecond") else v * 1e3)
        rows.append((r["Kernel Name"], r.get("Grid Size", ""), v))
tot = sum(v for _, _, v in rows)
agg = collections.OrderedDict()
for name, grid, v in rows:
    short = re.sub(r"\(.*$", "", name)
    short = re.sub(r"^void ", "", short)
    a = agg.setdefault(short, [0, 0.0])
    a[0] += 1
    a[1] += v
print(f"{len(rows)} launches, {tot:.1f} us total ({tot / per_step:.1f} us per step over {per_step} steps)")
print(f"{'kernel':70s} {'calls':>6s} {'us':>10s} {'share':>7s}")
for k, (n, v) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{k[:70]:70s} {n:6d} {v:10.1f} {100 * v / tot:6.1f}%")
if "--each" in sys.argv:
    for name, grid, v in rows:
        print(f"{v:9.1f} us  grid {grid:18s} {re.sub(r'^void ', '', name)[:110]}")
