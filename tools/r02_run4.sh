#!/bin/bash
# round 2, run 4: full GPU suite, the default bench line, ncu launch list of the same command and ncu --set full
# of the sweep kernel at the bench's launch shape
set -u
mkdir -p gpurun_out
T=${1:-r02h}
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/${T}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/${T}_pytest_gpu.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/${T}_bench_n1.json 2> gpurun_out/${T}_bench_n1.err; echo "bench rc=$?"; tail -3 gpurun_out/${T}_bench_n1.err
SMALL="python bench.py --steps 2 --warmup 3 --cards 74 --e2e-cards 16 --e2e-steps 1 --default-op-cards 32 --hough-batch 4 --no-cpu-baseline --no-oracle"
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 4000 --csv --log-file gpurun_out/${T}_launches.csv $SMALL > gpurun_out/${T}_ncu_launches.log 2>&1
echo "ncu launch list rc=$?"
QUICK="python bench.py --steps 1 --warmup 3 --e2e-cards 0 --no-configs --no-cpu-baseline --no-oracle"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:sweepw_kernel -s 16 -c 1 -f -o gpurun_out/${T}_sweepw $QUICK > gpurun_out/${T}_ncu_sweepw.log 2>&1
echo "ncu sweepw rc=$?"
python tools/launch_shares.py gpurun_out/${T}_launches.csv 2>/dev/null | head -40
