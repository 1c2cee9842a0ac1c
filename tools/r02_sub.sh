#!/bin/bash
set -u
mkdir -p gpurun_out
for B in 16 24 32; do
  OICS_FILES_SUBBATCH=$B timeout 900 python bench.py --steps 3 --warmup 3 --cards 64 --no-configs --no-cpu-baseline --no-oracle --e2e-steps 6 --e2e-cards ${1:-128} > gpurun_out/sb_$B.json 2> gpurun_out/sb_$B.err
  python - <<PY
import json
try:
    d=json.loads([l for l in open('gpurun_out/sb_$B.json') if l.startswith('{')][-1]); e=d['e2e']
    print('sub-batch $B: files', round(e['value'],1), 'raw', round(e['raw_pixels']['value'],1), 'jpeg_out', round(e['jpeg_out']['value'],1), 'cards', e['cards_per_gpu_per_step'])
except Exception as ex:
    print('failed', ex); print(open('gpurun_out/sb_$B.err').read()[-800:])
PY
done
