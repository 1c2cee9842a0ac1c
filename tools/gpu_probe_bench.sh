#!/bin/bash
# One gpurun call: the on-chip probe test and the default bench line (with roofline.on_chip).
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -k probes 2>&1 | tail -2
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
python - <<'P'
import json
d = json.loads(open("gpurun_out/bench.json").read().strip().splitlines()[-1])
print(d["value"], d["e2e"]["value"], json.dumps(d["roofline"]["on_chip"]))
P
