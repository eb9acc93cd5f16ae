"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel totals over the LAST n launches, or
(n = -1) over the last whole training step = everything after the optimizer launch that ended the previous step (the list
holds torch's own kernels too, so bench.py's `gpu_launches` per step is not the right count)."""
import csv
import re
import sys
from collections import defaultdict

path, last_n = sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 0
rows = []
with open(path, newline="") as f:
    lines = [ln for ln in f if ln.startswith('"')]
for r in csv.DictReader(lines):
    if r.get("Metric Name") != "gpu__time_duration.sum":
        continue
    v = float(r["Metric Value"].replace(",", ""))
    unit = r.get("Metric Unit", "ns")
    ns = v * {"ns": 1.0, "us": 1e3, "usecond": 1e3, "ms": 1e6, "msecond": 1e6, "nsecond": 1.0, "s": 1e9, "second": 1e9}.get(unit, 1.0)
    rows.append((re.sub(r"\(.*", "", r["Kernel Name"]), ns))
if last_n == -1:
    opt = [i for i, (n, _) in enumerate(rows) if "adam_dev_kernel" in n]  # two per step: D then G
    assert len(opt) >= 3, "need at least two steps in the list"
    rows = rows[opt[-3] + 1: opt[-1] + 1]
elif last_n:
    rows = rows[-last_n:]
agg = defaultdict(lambda: [0, 0.0])
for name, ns in rows:
    agg[name][0] += 1
    agg[name][1] += ns
total = sum(v[1] for v in agg.values())
print(f"launches {len(rows)}  total {total / 1e6:.2f} ms (cold-cache, serialised per-launch times)")
for name, (n, ns) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{ns / 1e6:9.3f} ms {100 * ns / total:5.1f}%  x{n:5d}  avg {ns / n / 1e3:9.1f} us  {name[:110]}")
