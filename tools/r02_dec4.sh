#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x -k "jpeg or files or decode or batch" 2>&1 | tail -2
python tools/codec_probe.py | python -c "
import sys,json
d=json.load(sys.stdin); print({k.replace('jdec_','').replace('_kernel',''):v for k,v in d['us_per_card'].items()})"
OICS_JD_WIDE=0 python tools/codec_probe.py | python -c "
import sys,json
d=json.load(sys.stdin); print('walk only', {k.replace('jdec_','').replace('_kernel',''):v for k,v in d['us_per_card'].items() if 'jdec' in k})"
timeout 900 python bench.py --steps 3 --warmup 3 --cards 64 --no-configs --no-cpu-baseline --e2e-steps 6 > gpurun_out/d4.json 2> gpurun_out/d4.err
python - <<'PY'
import json
try:
    d=json.loads([l for l in open('gpurun_out/d4.json') if l.startswith('{')][-1]); e=d['e2e']
    print('files', round(e['value'],1), 'raw', round(e['raw_pixels']['value'],1), 'jpeg_out', round(e['jpeg_out']['value'],1), 'not recovered', e['skew_not_recovered'], e.get('oracle'))
except Exception as ex:
    print('failed', ex); print(open('gpurun_out/d4.err').read()[-1500:])
PY
