#!/bin/bash
# e2e experiments: file path with different numbers of cards decoded concurrently
set -u
mkdir -p gpurun_out
TAG=${1:-e}
for B in 8 16 32; do
  OICS_FILES_SUBBATCH=$B timeout 600 python bench.py --steps 6 --warmup 3 --cards 64 --no-configs --no-cpu-baseline --no-oracle > gpurun_out/${TAG}_B$B.json 2> gpurun_out/${TAG}_B$B.err
  python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/${TAG}_B$B.json') if l.startswith('{')][-1])
e=d['e2e']
print('B=$B', 'files', round(e['value'],1), 'raw', round(e['raw_pixels']['value'],1), 'jpeg_out', round(e['jpeg_out']['value'],1))
PY
done
