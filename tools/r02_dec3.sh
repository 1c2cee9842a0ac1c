#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x -k "jpeg or files or decode" 2>&1 | tail -2
for WIDE in 128; do
  echo "== OICS_JD_WIDE=$WIDE"
  OICS_JD_WIDE=$WIDE python tools/codec_probe.py | python -c "
import sys,json
d=json.load(sys.stdin); print({k.replace('jdec_','').replace('_kernel',''):v for k,v in d['us_per_card'].items() if k.startswith('jdec')})"
  OICS_JD_WIDE=$WIDE timeout 900 python bench.py --steps 3 --warmup 3 --cards 64 --no-configs --no-cpu-baseline --no-oracle --e2e-steps 6 > gpurun_out/d3_$WIDE.json 2> gpurun_out/d3_$WIDE.err
  python - <<PY
import json
try:
    d=json.loads([l for l in open('gpurun_out/d3_$WIDE.json') if l.startswith('{')][-1]); e=d['e2e']
    print('value', round(d['value'],1), 'files', round(e['value'],1), 'raw', round(e['raw_pixels']['value'],1), 'jpeg_out', round(e['jpeg_out']['value'],1), 'not recovered', e['skew_not_recovered'])
except Exception as ex:
    print('failed', ex); print(open('gpurun_out/d3_$WIDE.err').read()[-1500:])
PY
done
bash tools/r02_timeline.sh > gpurun_out/d3_timeline.txt 2>&1; head -30 gpurun_out/d3_timeline.txt
