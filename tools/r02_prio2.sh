#!/bin/bash
set -u
mkdir -p gpurun_out
for P in 1000 0 -2; do
  OICS_LANE_PRIORITY=$P timeout 900 python bench.py --steps 3 --warmup 3 --cards 64 --no-configs --no-cpu-baseline --no-oracle --e2e-steps 6 > gpurun_out/pr_$P.json 2> gpurun_out/pr_$P.err
  python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/pr_$P.json') if l.startswith('{')][-1]); e=d['e2e']
print('lane priority $P: files', round(e['value'],1), 'raw', round(e['raw_pixels']['value'],1), 'jpeg_out', round(e['jpeg_out']['value'],1))
PY
done
