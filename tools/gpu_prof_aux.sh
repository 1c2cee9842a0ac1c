#!/bin/bash
# ncu --set full of the secondary (HBM-bound) kernels while configs 1-3 run on one A4 card.
set -u
mkdir -p gpurun_out
CMD="python tools/configs_bench.py"
timeout 600 $CMD > gpurun_out/configs.json 2> gpurun_out/configs.err; echo "plain rc=$?"
cat gpurun_out/configs.json
ncu --set full --clock-control none --import-source on -k regex:"${1:-warp_flat|gray_erode|canny_nms|fft_rows|fft_cols|fft_log|pack_bits|hough_accum}" -c ${2:-40} -o gpurun_out/prof_aux $CMD > gpurun_out/ncu_aux.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_aux.log
