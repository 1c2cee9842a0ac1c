#!/bin/bash
# GPU box: decoder tests + probe + launch list, then the bench line with e2e.jpeg_io
bash tools/gpu_jpegdec.sh ncu
python bench.py --no-cpu-baseline --steps 3 > gpurun_out/bench_jpeg.json 2> gpurun_out/bench_jpeg.err; echo "bench rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/bench_jpeg.json").read().strip().splitlines()[-1])
e = d["e2e"]
print("value", d["value"], "e2e", e["value"], "jpeg_out", e["jpeg_out"]["value"], "jpeg_io", e["jpeg_io"]["value"], e["jpeg_io"]["mismatches"])
PY
