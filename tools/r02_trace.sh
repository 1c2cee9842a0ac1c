#!/bin/bash
set -u
mkdir -p gpurun_out
OICS_BATCH_TRACE=1 timeout 900 python bench.py --steps 3 --warmup 3 --cards 64 --no-configs --no-cpu-baseline --no-oracle --e2e-steps 4 > gpurun_out/tr_bench.json 2> gpurun_out/tr_bench.err
grep "oics batch" gpurun_out/tr_bench.err | sed -n 41,56p | cut -c1-260
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/tr_bench.json') if l.startswith('{')][-1]); e=d['e2e']
print('value', round(d['value'],1), 'files', round(e['value'],1), 'raw', round(e['raw_pixels']['value'],1), 'jpeg_out', round(e['jpeg_out']['value'],1))
PY
