#!/bin/bash
set -u
mkdir -p gpurun_out
rm -f /tmp/tl.*.csv
OICS_PROFILE_TIMELINE=/tmp/tl timeout 900 python bench.py --steps 3 --warmup 3 --cards 64 --no-configs --no-cpu-baseline --no-oracle --e2e-steps 5 > gpurun_out/tl_bench.json 2> gpurun_out/tl_bench.err
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/tl_bench.json') if l.startswith('{')][-1]); e=d['e2e']
print('value', round(d['value'],1), 'files', round(e['value'],1), 'raw', round(e['raw_pixels']['value'],1), 'jpeg_out', round(e['jpeg_out']['value'],1))
PY
for f in /tmp/tl.*.csv; do
  if grep -q jdec_sync $f; then cp $f gpurun_out/tl_e2e.csv; python tools/timeline.py $f 60; fi
done
