#!/bin/bash
# Hough parity tests + PPHT timing on the A4 card (thr 0 and thr 100) + configs 1-3.
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "hough or fft_entry or correct" > gpurun_out/pytest_hough.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_hough.log
timeout 300 python tools/ppht_once.py; THR=100 timeout 300 python tools/ppht_once.py
timeout 600 python tools/configs_bench.py > gpurun_out/configs.json 2> gpurun_out/configs.err; echo "configs rc=$?"; cat gpurun_out/configs.json; tail -3 gpurun_out/configs.err
