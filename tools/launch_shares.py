#!/usr/bin/env python
"""Per-kernel totals of an `ncu --csv` launch list (gpu__time_duration and, if present, DRAM bytes)."""
import collections, csv, sys
rows = list(csv.reader(open(sys.argv[1])))
for i, r in enumerate(rows):
    if r and r[0] == 'ID':
        hdr, start = r, i + 1
        break
iN, iM, iV, iU = hdr.index('Kernel Name'), hdr.index('Metric Name'), hdr.index('Metric Value'), hdr.index('Metric Unit')
t = collections.defaultdict(lambda: collections.Counter())
n = collections.Counter()
scale = {'ns': 1e-3, 'us': 1.0, 'ms': 1e3, 'byte': 1e-6, 'Kbyte': 1e-3, 'Mbyte': 1.0, 'Gbyte': 1e3}
for r in rows[start:]:
    if len(r) <= iV:
        continue
    name = r[iN].split('(')[0]
    t[name][r[iM]] += float(r[iV].replace(',', '')) * scale.get(r[iU], 1.0)
    if r[iM] == 'gpu__time_duration.sum':
        n[name] += 1
tot = sum(v['gpu__time_duration.sum'] for v in t.values())
for k, v in sorted(t.items(), key=lambda kv: -kv[1]['gpu__time_duration.sum']):
    us = v['gpu__time_duration.sum']
    mb = v.get('dram__bytes_read.sum', 0) + v.get('dram__bytes_write.sum', 0)
    cyc = v.get('sm__cycles_elapsed.max', 0)
    ghz = f" {cyc/us*1e-3:5.2f} GHz" if cyc else ""
    print(f"{k:44s} n={n[k]:4d}{ghz} total {us/1e3:8.3f} ms {100*us/tot:5.1f}%  per launch {us/n[k]:8.1f} us  dram {mb/n[k]:8.1f} MB/launch  {mb/us*1e-3 if us else 0:6.2f} TB/s")
