#!/bin/bash
# GPU box: decoder parity tests, the decoder probe and its ncu launch list.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu -x -k "jpeg or correct_batch" > gpurun_out/jpegdec_tests.log 2>&1; echo "tests rc=$?"
tail -15 gpurun_out/jpegdec_tests.log
timeout 300 python tools/jpegdec_probe.py > gpurun_out/jpegdec_probe.json 2> gpurun_out/jpegdec_probe.err; echo "probe rc=$?"; cat gpurun_out/jpegdec_probe.json; tail -3 gpurun_out/jpegdec_probe.err
if [ "$1" = "ncu" ]; then
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_jpegdec.csv \
    python tools/jpegdec_probe.py > gpurun_out/ncu_jpegdec.log 2>&1; echo "ncu rc=$?"
python tools/launch_shares.py gpurun_out/launches_jpegdec.csv 2>/dev/null | grep -i "jdec\|scan"
fi
