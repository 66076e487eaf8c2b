#!/usr/bin/env python
"""per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv` launch list:
    python profiles/launch_summary.py profiles/r1_launches_final.csv "<command line that was profiled>" > summary.md"""
import collections
import csv
import re
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
hdr, agg = None, collections.OrderedDict()
for r in rows:
    if r[0] == 'ID':
        hdr = r
        continue
    if hdr is None:
        continue
    d = dict(zip(hdr, r))
    if d.get('Metric Name') != 'gpu__time_duration.sum':
        continue
    name = re.sub(r'\(.*', '', d['Kernel Name']).replace('void ', '')
    v = float(d['Metric Value'].replace(',', ''))
    u = d['Metric Unit']
    if u in ('nsecond', 'ns'):
        v /= 1e3
    elif u in ('msecond', 'ms'):
        v *= 1e3
    a = agg.setdefault(name, [0, 0.0])
    a[0] += 1
    a[1] += v
tot = sum(a[1] for a in agg.values())
print("# ncu launch list (gpu__time_duration.sum, --clock-control none)\n")
print("Command: `%s` (cold-cache, serialised: compare shares).\n" % (sys.argv[2] if len(sys.argv) > 2 else "?"))
print("| kernel | launches | total us | avg us | share |\n|---|---|---|---|---|")
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print("| %s | %d | %.1f | %.1f | %.3f |" % (k, a[0], a[1], a[1] / a[0], a[1] / tot))


def get(prefix):
    for k, a in agg.items():
        if k.startswith(prefix):
            return a
    return None


three = [(n, get(n)) for n in ('kstar_mean_i8_kernel', 'i8::colsumsq_i8_kernel', 'acquire_epilogue_kernel')]
if all(a for _, a in three):
    t3 = sum(a[1] / a[0] for _, a in three)
    print("\nPer candidate batch of the acquisition step, serialised under ncu:\n")
    for n, a in three:
        print("* %s: %.1f us avg (%.1f %% of the three)" % (n, a[1] / a[0], 100 * a[1] / a[0] / t3))
    print("\nThe bench's live CUDA-event measurement of the variance kernel inside the two-stream pipeline gives its share "
          "of the\nstep as `roofline.gemm_share_of_step`; under ncu the three kernels of a batch run one after the other, "
          "so the\nvariance kernel's share of the three (above) is the comparable figure.")
