#!/bin/bash
# ncu --set full of the secondary kernels (regex in $1) in the small bench
set -u
mkdir -p gpurun_out
SMALL="python bench.py --steps 1 --warmup 3 --cards 16 --e2e-cards 8 --no-cpu-baseline"
$SMALL > gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"$1" -s ${2:-20} -c ${3:-3} -o gpurun_out/prof_aux $SMALL > gpurun_out/ncu3.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu3.log
