#!/bin/bash
set -u
mkdir -p gpurun_out
OICS_BATCH_TRACE=1 timeout 900 python bench.py --steps 3 --warmup 3 --cards 64 --no-configs --no-cpu-baseline --no-oracle --e2e-steps 6 > gpurun_out/pp.json 2> gpurun_out/pp.err
grep "oics batch" gpurun_out/pp.err | grep -v "begin -\|finish -" | sed -n 19,24p | cut -c1-250
bash tools/r02_timeline.sh > gpurun_out/pp_timeline.txt 2>&1; sed -n 1,56p gpurun_out/pp_timeline.txt
