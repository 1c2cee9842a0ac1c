#!/bin/bash
set -u
mkdir -p gpurun_out
TAG=${1:-d}
python tools/profile_decode.py | python -c "
import sys,json
d=json.load(sys.stdin); print({k['kernel'].replace('jdec_','').replace('_kernel',''):k['us'] for k in d['kernels']})"
timeout 600 python -m pytest tests -m gpu -q -x -k "jpeg" 2>&1 | tail -2
timeout 600 python bench.py --steps 6 --warmup 3 --cards 64 --no-configs --no-cpu-baseline --no-oracle > gpurun_out/${TAG}.json 2> gpurun_out/${TAG}.err
python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/${TAG}.json') if l.startswith('{')][-1])
e=d['e2e']
print('files', round(e['value'],1), 'raw', round(e['raw_pixels']['value'],1), 'jpeg_out', round(e['jpeg_out']['value'],1), 'not recovered', e['skew_not_recovered'])
PY
