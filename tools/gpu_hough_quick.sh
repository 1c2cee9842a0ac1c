#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "hough" > gpurun_out/pytest_hough.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_hough.log
timeout 300 python tools/ppht_once.py; THR=100 timeout 300 python tools/ppht_once.py
