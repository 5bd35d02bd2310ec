"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list into per-kernel time of ONE train step.
python tools/launch_summary.py launches.csv [step_index]   (steps are delimited by clip_transform_kernel / tubelet_transform_kernel, one per step)"""
import collections
import csv
import re
import sys

rows = []
with open(sys.argv[1]) as f:
    lines = [l for l in f if not l.startswith("==")]
rd = csv.DictReader(lines)
for r in rd:
    if r.get("Metric Name") != "gpu__time_duration.sum":
        continue
    v = float(r["Metric Value"].replace(",", ""))
    unit = r.get("Metric Unit", "ns")
    ns = v * {"ns": 1.0, "us": 1e3, "ms": 1e6, "s": 1e9}.get(unit, 1.0)
    rows.append((r["Kernel Name"], ns))
starts = [i for i, (k, _) in enumerate(rows) if "clip_transform_kernel" in k or "tubelet_transform_kernel" in k]
which = int(sys.argv[2]) if len(sys.argv) > 2 else len(starts) - 2
lo, hi = starts[which], (starts[which + 1] if which + 1 < len(starts) else len(rows))
step = rows[lo:hi]
agg = collections.OrderedDict()
for k, ns in step:
    k = re.sub(r"^void ", "", k)
    k = re.sub(r"<unnamed>::", "", k)
    k = re.sub(r"\(.*$", "", k)
    c, t = agg.get(k, (0, 0.0))
    agg[k] = (c + 1, t + ns)
tot = sum(t for _, t in agg.values())
print(f"train step #{which} of {len(starts)} captured: {len(step)} launches, {tot / 1e6:.1f} ms of serialised cold-cache kernel time")
print(f"{'kernel':64s} {'count':>5s} {'ms':>9s}  share")
for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{k[:64]:64s} {c:5d} {t / 1e6:9.2f} {100 * t / tot:5.1f}%")
