#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x -k "jpeg or files or decode or batch" 2>&1 | tail -2
for L in 0 2 3 4; do
  OICS_JPEG_LANES_N=$L timeout 900 python bench.py --steps 3 --warmup 3 --cards 64 --no-configs --no-cpu-baseline --e2e-steps 6 > gpurun_out/ln_$L.json 2> gpurun_out/ln_$L.err
  python - <<PY
import json
try:
    d=json.loads([l for l in open('gpurun_out/ln_$L.json') if l.startswith('{')][-1]); e=d['e2e']
    print('lanes $L: files', round(e['value'],1), 'raw', round(e['raw_pixels']['value'],1), 'jpeg_out', round(e['jpeg_out']['value'],1), 'not recovered', e['skew_not_recovered'], e.get('oracle'))
except Exception as ex:
    print('lanes $L failed', ex); print(open('gpurun_out/ln_$L.err').read()[-1200:])
PY
done
