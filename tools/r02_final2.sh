#!/bin/bash
# end-of-round confirmation on one B200: whole GPU suite, the acceptance run on the fixtures, the bench line of both arms,
# ncu of the sweep kernel, the launch list of a small bench, the codec kernels alone
set -u
mkdir -p gpurun_out
T=${1:-r02zz}
timeout 1500 python -m pytest tests -m gpu -q -x 2>&1 | tail -3
timeout 1500 python tools/accuracy_regression.py > gpurun_out/${T}_accuracy_regression_dataset.json 2> gpurun_out/${T}_accuracy.err; echo "accuracy rc=$?"
python - <<PY
import json
d=json.load(open('gpurun_out/${T}_accuracy_regression_dataset.json'))
print({k:d[k] for k in ('histogram','inputs','average_error_deg','gpu_cards_per_s','oracle_checked')}, 'mismatches', len(d['oracle_mismatches']))
PY
bash tools/r02_final.sh $T
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${T}_launches_bench_small.csv python bench.py --steps 1 --warmup 3 --cards 64 --e2e-cards 32 --e2e-steps 1 --no-configs --no-cpu-baseline --no-oracle > gpurun_out/${T}_ncu_launches.log 2>&1; echo "launch list rc=$?"
python tools/codec_probe.py > gpurun_out/${T}_codec_kernels.json; cat gpurun_out/${T}_codec_kernels.json
