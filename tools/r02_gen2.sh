#!/bin/bash
set -u
mkdir -p gpurun_out
python tools/sweep_general_probe.py > gpurun_out/r02g3_probe.json 2> gpurun_out/r02g3_probe.err; cat gpurun_out/r02g3_probe.json; tail -2 gpurun_out/r02g3_probe.err
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "sweep or accuracy or correct_default or entry or two_contexts or 65535" 2>&1 | tail -2
timeout 600 python bench.py --steps 3 --warmup 3 --cards 64 --e2e-cards 0 --no-cpu-baseline --hough-batch 0 --default-op-cards 32 2>/dev/null | python -c "
import sys,json
d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); c=d['configs']['c5_huge_scan']; print('c5', c['seconds'], c['oracle_equal'])"
