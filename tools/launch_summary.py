#!/usr/bin/env python
"""Aggregate an ncu launch list (--metrics gpu__time_duration.sum --csv) per kernel: launches, total / average ms, share.
usage: python tools/launch_summary.py profiles/r01_launches.csv > profiles/r01_launches_summary.txt"""
import csv, sys
from collections import OrderedDict

rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if l.startswith('"'))]
hdr = rows[0]
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
agg = OrderedDict()
for r in rows[1:]:
    v = float(r[vi].replace(",", ""))
    ms = v / 1e6 if r[ui] in ("ns", "nsecond") else v / 1e3 if r[ui] in ("us", "usecond") else v
    a = agg.setdefault(r[ki], [0, 0.0])
    a[0] += 1; a[1] += ms
total = sum(a[1] for a in agg.values())
print(f"# {sum(a[0] for a in agg.values())} launches captured, total {total:.2f} ms (cold-cache, serialised: compare SHARES with bench.py's live CUDA-event numbers)")
steps = {k: a for k, a in agg.items() if "msm_" in k and "precompute" not in k}
st = sum(a[1] for a in steps.values())
acc = sum(a[1] for k, a in steps.items() if "msm_accumulate" in k)
if st:
    print(f"# within the captured MSM steps: accumulate {acc:.1f} ms of {st:.1f} ms = {acc / st:.3f}")
print(f"{'kernel':80s} {'launches':>8s} {'total_ms':>10s} {'avg_ms':>10s} {'share':>7s}")
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{k[:80]:80s} {a[0]:8d} {a[1]:10.3f} {a[1] / a[0]:10.4f} {a[1] / total:7.3f}")
