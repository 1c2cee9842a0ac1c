#!/bin/bash
# ncu launch list (durations + DRAM bytes) of every kernel of configs 1-3 on one A4 card.
set -u
mkdir -p gpurun_out
timeout 600 python tools/configs_bench.py > gpurun_out/configs.json 2> gpurun_out/configs.err && \
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__cycles_elapsed.max --clock-control none -c 3000 --csv --log-file gpurun_out/launches_configs.csv python tools/configs_bench.py > gpurun_out/ncu_cfg.log 2>&1
echo "ncu rc=$?"
python tools/launch_shares.py gpurun_out/launches_configs.csv
python tools/kernel_rooflines.py gpurun_out/launches_configs.csv gpurun_out/kernel_rooflines.json
