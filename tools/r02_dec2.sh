#!/bin/bash
set -u
mkdir -p gpurun_out
python tools/codec_probe.py
timeout 900 python -m pytest tests -m gpu -q -x -k "jpeg or files or decode" 2>&1 | tail -2
for MC in 8 32; do
  CUDA_DEVICE_MAX_CONNECTIONS=$MC timeout 900 python bench.py --steps 3 --warmup 3 --cards 64 --no-configs --no-cpu-baseline --no-oracle --e2e-steps 6 > gpurun_out/d2_$MC.json 2> gpurun_out/d2_$MC.err
  python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/d2_$MC.json') if l.startswith('{')][-1]); e=d['e2e']
print('connections $MC: value', round(d['value'],1), 'files', round(e['value'],1), 'raw', round(e['raw_pixels']['value'],1), 'jpeg_out', round(e['jpeg_out']['value'],1), 'not recovered', e['skew_not_recovered'])
PY
done
bash tools/r02_timeline.sh | head -24
