#!/bin/bash
set -u
mkdir -p gpurun_out
TAG=${1:-e}
for T in 1 2 3; do
  timeout 600 python bench.py --steps 6 --warmup 3 --cards 64 --e2e-contexts $T --no-configs --no-cpu-baseline --no-oracle > gpurun_out/${TAG}_T$T.json 2> gpurun_out/${TAG}_T$T.err
  python - <<PY
import json
try:
    d=json.loads([l for l in open('gpurun_out/${TAG}_T$T.json') if l.startswith('{')][-1])
    e=d['e2e']
    print('T=$T', 'files', round(e['value'],1), 'raw', round(e['raw_pixels']['value'],1), 'jpeg_out', round(e['jpeg_out']['value'],1), 'not recovered', e['skew_not_recovered'])
except Exception as ex:
    print('T=$T failed', ex); print(open('gpurun_out/${TAG}_T$T.err').read()[-1500:])
PY
done
