#!/usr/bin/env python
"""Read an OICS_PROFILE_TIMELINE dump ("kernel",start_us,end_us per launch) and print, per kernel, launches /
summed duration / first start / last end, then the step in bins: which kernel families were running."""
import csv, sys, collections
rows = [(r[0], float(r[1]), float(r[2])) for r in csv.reader(open(sys.argv[1]))]
nb = int(sys.argv[2]) if len(sys.argv) > 2 else 60
t_end = max(r[2] for r in rows)
print(f"{len(rows)} launches, span {t_end / 1000:.2f} ms")
agg = collections.OrderedDict()
for n, a, b in rows:
    e = agg.setdefault(n, [0, 0.0, a, b])
    e[0] += 1; e[1] += b - a; e[2] = min(e[2], a); e[3] = max(e[3], b)
for n, (c, d, a, b) in agg.items():
    print(f"  {n:34s} {c:5d} launches  sum {d / 1000:8.2f} ms  first {a / 1000:7.2f}  last end {b / 1000:7.2f}")
def fam(n):
    if n.startswith('jdec_sync'): return 'S'
    if n.startswith('jdec_write'): return 'W'
    if n.startswith('jdec'): return 'D'
    if n.startswith('jpeg'): return 'E'
    if n.startswith('sweepw_kernel') or n.startswith('sweep_kernel') or n.startswith('sweeps'): return 'X'
    if n.startswith('warp'): return 'R'
    return 'o'
fams = 'XSWDERo'
bw = t_end / nb
print("bins of %.2f ms; per family the summed kernel-time in the bin / bin width (= kernels of that family in flight)" % (bw / 1000))
print("  bin   " + "".join(f"{f:>7s}" for f in fams))
for i in range(nb):
    lo, hi = i * bw, (i + 1) * bw
    acc = collections.Counter()
    for n, a, b in rows:
        ov = min(b, hi) - max(a, lo)
        if ov > 0: acc[fam(n)] += ov
    print(f"  {lo / 1000:6.1f} " + "".join(f"{acc[f] / bw:7.2f}" for f in fams))
