#!/bin/bash
# round 2, run 1: all GPU parity tests, the full default bench line, ncu --set full of the sweep
# kernel at the bench's launch shape (36 cards x 300 angles) and of the general kernel on config 5.
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/r02a_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/r02a_pytest_gpu.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/r02a_bench_n1.json 2> gpurun_out/r02a_bench_n1.err; echo "bench rc=$?"; tail -3 gpurun_out/r02a_bench_n1.err
QUICK="python bench.py --steps 1 --warmup 3 --e2e-cards 0 --no-configs --no-cpu-baseline --no-oracle"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:sweepw_kernel -s 16 -c 1 -f -o gpurun_out/r02a_sweepw $QUICK > gpurun_out/r02a_ncu_sweepw.log 2>&1
echo "ncu sweepw rc=$?"; tail -2 gpurun_out/r02a_ncu_sweepw.log
C5="python bench.py --steps 1 --warmup 3 --cards 16 --e2e-cards 0 --no-cpu-baseline --no-oracle"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"sweep_kernel" -s 2 -c 1 -f -o gpurun_out/r02a_sweepg $C5 > gpurun_out/r02a_ncu_sweepg.log 2>&1
echo "ncu sweep general rc=$?"; tail -2 gpurun_out/r02a_ncu_sweepg.log
ls -la gpurun_out | tail -12
