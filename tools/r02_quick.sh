#!/bin/bash
# quick loop: the sweep parity tests + the headline number (no e2e, no configs)
set -u
mkdir -p gpurun_out
TAG=${1:-q}
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "${2:-sweep or correct or hough or edges or accuracy}" > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/${TAG}_pytest.log
timeout 600 python bench.py --steps 10 --warmup 3 --e2e-cards 0 --no-configs --no-cpu-baseline > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/${TAG}_bench.err
python - <<PY
import json
d=json.loads(open('gpurun_out/${TAG}_bench.json').read().strip().splitlines()[-1])
print('value', d['value'], 'frac', d['roofline']['frac'], 'launch ms', d['roofline']['avg_launch_ms'], 'parity', d['parity'].get('oracle'))
PY
