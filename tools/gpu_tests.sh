#!/bin/bash
# GPU parity tests (all) + ncu launch list of a small bench (kernel durations of every kernel).
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
SMALL="python bench.py --steps 2 --warmup 3 --cards 16 --e2e-cards 8 --no-cpu-baseline"
$SMALL > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv $SMALL > gpurun_out/ncu1.log 2>&1
echo "ncu rc=$?"
python tools/launch_shares.py gpurun_out/launches.csv
