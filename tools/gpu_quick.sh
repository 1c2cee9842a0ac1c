#!/bin/bash
# Quick GPU iteration: sweep parity tests + bench (no CPU baseline).
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -k "sweep or entry or correct" > gpurun_out/pytest_quick.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/pytest_quick.log
python bench.py --no-cpu-baseline --e2e-cards 128 "$@" > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench rc=$?"
tail -3 gpurun_out/bench_quick.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_quick.json'))
print({k:d[k] for k in ('value','ms_per_step','gpu_launches','parity')}, d['roofline']['frac'], d['roofline']['avg_launch_ms'], d['roofline']['kernel_share_of_step'], d['e2e'])
PY
