#!/bin/bash
# GPU box: full ncu captures of the decoder's two entropy kernels on the A4 card's file
# (jdec_sync_kernel: the first three launches = round 0, the plain step, the first chain round).
mkdir -p gpurun_out
timeout 300 python tools/jpegdec_probe.py > gpurun_out/jpegdec_probe.json 2> gpurun_out/jpegdec_probe.err; echo "probe rc=$?"; cat gpurun_out/jpegdec_probe.json
timeout 600 ncu --set full --clock-control none --import-source on -k regex:jdec_write_kernel -s 1 -c 1 -f -o gpurun_out/full_jdec_write_kernel \
    python tools/jpegdec_probe.py > gpurun_out/ncu_full_jdec_write_kernel.log 2>&1; echo "ncu write rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:jdec_sync_kernel -s 0 -c 3 -f -o gpurun_out/full_jdec_sync_kernel \
    python tools/jpegdec_probe.py > gpurun_out/ncu_full_jdec_sync_kernel.log 2>&1; echo "ncu sync rc=$?"
