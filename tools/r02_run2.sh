#!/bin/bash
# round 2, run 2: GPU tests, bench, the full N4 acceptance run on the reference's dataset, compute-sanitizer logs
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/r02b_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/r02b_pytest_gpu.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/r02b_bench_n1.json 2> gpurun_out/r02b_bench_n1.err; echo "bench rc=$?"; tail -3 gpurun_out/r02b_bench_n1.err
timeout 1200 python tools/accuracy_regression.py > gpurun_out/r02b_accuracy_regression_dataset.json 2> gpurun_out/r02b_accuracy_regression.err; echo "N4 rc=$?"; tail -2 gpurun_out/r02b_accuracy_regression.err
for tool in memcheck racecheck synccheck; do
  timeout 900 compute-sanitizer --tool $tool --print-limit 20 python tools/sanitize_small.py > gpurun_out/r02b_sanitizer_$tool.log 2>&1
  echo "sanitizer $tool rc=$?"; tail -4 gpurun_out/r02b_sanitizer_$tool.log
done
