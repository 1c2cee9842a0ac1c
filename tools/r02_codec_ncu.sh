#!/bin/bash
# per-line ncu of the codec kernels named on the command line (default: the full-occupancy ones)
set -u
mkdir -p gpurun_out
python tools/codec_probe.py
for K in ${@:-jdec_idct_kernel jdec_color_kernel jpeg_dct_kernel jpeg_emit_kernel}; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:"$K" -s 2 -c 1 -f -o /tmp/cod_$K python tools/codec_probe.py 1 > gpurun_out/cod_$K.log 2>&1; echo "== $K ncu rc=$?"
  python tools/ncu_summary.py /tmp/cod_$K.ncu-rep gpurun_out/r02c_${K}_summary.json > /dev/null 2>&1
  python - <<PY
import json
d=json.load(open('gpurun_out/r02c_${K}_summary.json'))
k=d[0] if isinstance(d,list) else d
k=k.get('counters',k) if isinstance(k,dict) else k
print({a:b for a,b in (k.items() if isinstance(k,dict) else []) if any(s in a for s in ('duration','registers','issue_active','warps_active','pipe_alu','pipe_lsu','inst_executed.sum','dram__bytes'))})
PY
  python tools/ncu_lines.py /tmp/cod_$K.ncu-rep 22
done
