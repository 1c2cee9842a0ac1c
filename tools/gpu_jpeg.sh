#!/bin/bash
# GPU box: JPEG parity tests, the encoder probe, its ncu launch list, and the bench with e2e.jpeg_out.
mkdir -p gpurun_out
python -m pytest tests -q -m gpu -x -k "jpeg or correct_batch" > gpurun_out/jpeg_tests.log 2>&1; echo "tests rc=$?"
tail -3 gpurun_out/jpeg_tests.log
python tools/jpeg_probe.py > gpurun_out/jpeg_probe.json 2> gpurun_out/jpeg_probe.err; echo "probe rc=$?"; cat gpurun_out/jpeg_probe.json; tail -3 gpurun_out/jpeg_probe.err
python bench.py --no-cpu-baseline --steps 3 > gpurun_out/bench_jpeg.json 2> gpurun_out/bench_jpeg.err; echo "bench rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/bench_jpeg.json").read().strip().splitlines()[-1])
print(d["value"], d["e2e"])
PY
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_jpeg.csv \
    python tools/jpeg_probe.py > gpurun_out/ncu_jpeg.log 2>&1; echo "ncu rc=$?"
python tools/launch_shares.py gpurun_out/launches_jpeg.csv 2>/dev/null | grep -i "jpeg\|scan" 
