#!/bin/bash
# ncu --set full summaries (+ per-line shares) of the codec kernels of the final build, one A4 card (tools/codec_probe.py 1:
# three decodes + encodes; the capture takes a launch of the second)
set -u
mkdir -p gpurun_out
cap() {  # kernel regex, launches to skip, tag
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:"$1" -s $2 -c 1 -f -o /tmp/cod_$3 python tools/codec_probe.py 1 > gpurun_out/cod_$3.log 2>&1; echo "== $3 ncu rc=$?"
  python tools/ncu_summary.py /tmp/cod_$3.ncu-rep gpurun_out/r02zz_$3_summary.json > /dev/null 2>&1
  python tools/ncu_lines.py /tmp/cod_$3.ncu-rep 14 > gpurun_out/r02zz_$3_lines.txt 2>&1
  python - <<PY
import json
for k in json.load(open('gpurun_out/r02zz_$3_summary.json')):
    print({a:b for a,b in k.items() if any(s in a for s in ('duration','registers','issue_active','warps_active','inst_executed.sum','dram__bytes'))})
PY
}
cap jdec_sync_kernel 3 jdec_sync_wide
cap jdec_walk_kernel 3 jdec_walk
cap jdec_idct_kernel 1 jdec_idct
cap jdec_color420_kernel 1 jdec_color420
cap jpeg_dct_kernel 1 jpeg_dct
cap jpeg_emit_kernel 1 jpeg_emit
