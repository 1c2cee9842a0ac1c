#!/bin/bash
set -u
mkdir -p gpurun_out
T=${1:-r02q}
timeout 1200 python tools/accuracy_regression.py --cards ${2:-16} --oracle-every 0 > gpurun_out/${T}_n4.json 2> gpurun_out/${T}_n4.err; echo "N4 rc=$?"; tail -1 gpurun_out/${T}_n4.err
python -c "
import json; d=json.load(open('gpurun_out/${T}_n4.json')); print(d['inputs'], d['histogram'], round(d['gpu_cards_per_s'],1))"
