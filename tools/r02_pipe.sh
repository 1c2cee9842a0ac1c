#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x -k "jpeg or files or decode or batch" 2>&1 | tail -2
OICS_BATCH_TRACE=1 timeout 900 python bench.py --steps 3 --warmup 3 --cards 64 --no-configs --no-cpu-baseline --no-oracle --e2e-steps 6 > gpurun_out/pp.json 2> gpurun_out/pp.err
grep "oics batch" gpurun_out/pp.err | grep -v "begin -\|finish -" | sed -n 19,24p | cut -c1-250
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/pp.json') if l.startswith('{')][-1]); e=d['e2e']
print('value', round(d['value'],1), 'files', round(e['value'],1), 'raw', round(e['raw_pixels']['value'],1), 'jpeg_out', round(e['jpeg_out']['value'],1), 'not recovered', e['skew_not_recovered'], 'oracle', e.get('oracle'))
PY
bash tools/r02_timeline.sh > gpurun_out/pp_timeline.txt 2>&1; sed -n 1,50p gpurun_out/pp_timeline.txt
