#!/bin/bash
set -u
mkdir -p gpurun_out
for P in 0 -1 -3; do
  OICS_BENCH_STREAM_PRIORITY=$P timeout 600 python bench.py --steps 6 --warmup 3 --cards 64 --no-configs --no-cpu-baseline --no-oracle > gpurun_out/r02w_P$P.json 2> gpurun_out/r02w_P$P.err
  python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/r02w_P$P.json') if l.startswith('{')][-1])
e=d['e2e']
print('prio=$P', 'files', round(e['value'],1), 'raw', round(e['raw_pixels']['value'],1), 'jpeg_out', round(e['jpeg_out']['value'],1), 'value', round(d['value'],1))
PY
done
