#!/bin/bash
set -u
mkdir -p gpurun_out
export OICS_JD_WIDE=0
K=jdec_walk_kernel
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"$K" -s 6 -c 1 -f -o /tmp/cod_$K python tools/codec_probe.py 1 > gpurun_out/cod_$K.log 2>&1; echo "== $K ncu rc=$?"
python tools/ncu_summary.py /tmp/cod_$K.ncu-rep gpurun_out/r02d_${K}_summary.json > /dev/null 2>&1
python - <<PY
import json
d=json.load(open('gpurun_out/r02d_${K}_summary.json'))
for k in d: print({a:b for a,b in k.items() if any(s in a for s in ('Name','Grid','duration','registers','issue_active','warps_active','inst_executed.sum','cycles_elapsed','opcode'))})
PY
python tools/ncu_lines.py /tmp/cod_$K.ncu-rep 45
