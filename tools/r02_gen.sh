#!/bin/bash
set -u
mkdir -p gpurun_out
python tools/sweep_general_probe.py > gpurun_out/r02g2_probe.json 2> gpurun_out/r02g2_probe.err; cat gpurun_out/r02g2_probe.json; tail -2 gpurun_out/r02g2_probe.err
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"sweep_kernel" -s 1 -c 1 -f -o /tmp/r02g2 python tools/sweep_general_probe.py > gpurun_out/r02g2_ncu.log 2>&1; echo "ncu rc=$?"
python tools/ncu_summary.py /tmp/r02g2.ncu-rep gpurun_out/r02g2_sweep_general_summary.json > /dev/null 2>&1
python tools/ncu_lines.py /tmp/r02g2.ncu-rep 30
