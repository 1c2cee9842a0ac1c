#!/bin/bash
# GPU box: encoder probe (parity asserted inside), launch list, one full capture of the kernel named in $1.
K=${1:-jpeg_dct_kernel}
mkdir -p gpurun_out
python tools/jpeg_probe.py > gpurun_out/jpeg_probe.json 2> gpurun_out/jpeg_probe.err; echo "probe rc=$?"; cat gpurun_out/jpeg_probe.json; tail -3 gpurun_out/jpeg_probe.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_jpeg.csv \
    python tools/jpeg_probe.py > gpurun_out/ncu_jpeg.log 2>&1; echo "ncu rc=$?"
python tools/launch_shares.py gpurun_out/launches_jpeg.csv 2>/dev/null | grep -i "jpeg\|scan"
ncu --set full --clock-control none --import-source on -k regex:$K -s 2 -c 1 -f -o gpurun_out/full_$K \
    python tools/jpeg_probe.py > gpurun_out/ncu_full_jpeg.log 2>&1; echo "ncu full rc=$?"
