#!/usr/bin/env python
"""Aggregate an ncu launch list (gpu__time_duration.sum per launch, --csv) into per-kernel totals for the LAST optimizer
step in the capture (delimited by the optimizer-step kernels). usage: launch_breakdown.py launches.csv [out.md]"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
h = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
hdr = rows[h]
kn, mv = hdr.index("Kernel Name"), hdr.index("Metric Value")
ev = []
for r in rows[h + 1:]:
    if len(r) > mv:
        try:
            ev.append((r[kn], float(r[mv].replace(",", ""))))
        except ValueError:
            pass
ends = [i for i, (k, _) in enumerate(ev) if "zero_step" in k or "adamw" in k.lower()]
# one optimizer step = launches after the previous step's last optimizer kernel up to this step's last one
groups = []
for i in ends:
    if groups and i - groups[-1][-1] < 8:
        groups[-1].append(i)
    else:
        groups.append([i])
if len(groups) >= 2:
    lo, hi = groups[-2][-1] + 1, groups[-1][-1] + 1
else:
    lo, hi = 0, len(ev)
step = ev[lo:hi]
agg = collections.defaultdict(lambda: [0, 0.0])
for k, v in step:
    k = k.replace("void ", "").replace("b200::", "")
    k = k.split("(")[0][:100]
    agg[k][0] += 1
    agg[k][1] += v
tot = sum(v[1] for v in agg.values())
out = [f"launches in the last optimizer step: {len(step)}; summed kernel time {tot / 1e6:.3f} ms", "",
       "| % of GPU time | launches | mean µs | kernel |", "|---:|---:|---:|---|"]
for k, v in sorted(agg.items(), key=lambda x: -x[1][1]):
    out.append(f"| {v[1] / tot * 100:.1f} | {v[0]} | {v[1] / v[0] / 1000:.1f} | `{k}` |")
txt = "\n".join(out)
print(txt)
if len(sys.argv) > 2:
    open(sys.argv[2], "w").write(txt + "\n")
