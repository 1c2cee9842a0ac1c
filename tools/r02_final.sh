#!/bin/bash
# final single-GPU line of the round + ncu --set full of the sweep kernel (report kept outside gpurun_out)
set -u
mkdir -p gpurun_out
T=${1:-r02u}
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/${T}_bench_n1.json 2> gpurun_out/${T}_bench_n1.err; echo "bench rc=$?"; tail -3 gpurun_out/${T}_bench_n1.err
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/${T}_bench_reference_n1.json 2> gpurun_out/${T}_bench_reference_n1.err; echo "reference rc=$?"
QUICK="python bench.py --steps 1 --warmup 3 --e2e-cards 0 --no-configs --no-cpu-baseline --no-oracle"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:sweepw_kernel -s 16 -c 1 -f -o /tmp/${T}_sweepw $QUICK > gpurun_out/${T}_ncu_sweepw.log 2>&1
echo "ncu sweepw rc=$?"
python tools/ncu_summary.py /tmp/${T}_sweepw.ncu-rep gpurun_out/${T}_sweepw_kernel_summary.json > /dev/null 2>&1
python tools/ncu_lines.py /tmp/${T}_sweepw.ncu-rep 30 > gpurun_out/${T}_sweepw_kernel_lines.txt 2>&1
python - <<PY
import json
b=json.loads([l for l in open('gpurun_out/${T}_bench_n1.json') if l.startswith('{')][-1])
r=b['roofline']
print('value',round(b['value'],1),'ms/step',round(b['ms_per_step'],2),'frac',round(r['frac'],4),'peak',round(r['peak']),'launch ms',round(r['avg_launch_ms'],3))
e=b['e2e']; print('e2e',round(e['value'],1),'raw',round(e['raw_pixels']['value'],1),'jpeg_out',round(e['jpeg_out']['value'],1), e['oracle'])
c=b['configs']
print('c2',round(c['c2_hough']['seconds'],4),c['c2_hough']['oracle_equal'],c['c2_hough']['batch']['cards_per_s'],'c3',round(c['c3_fft']['seconds'],4),c['c3_fft']['oracle_equal'],'c5',round(c['c5_huge_scan']['seconds'],4),c['c5_huge_scan']['oracle_equal'])
d=c['default_op']; print('dop host',round(d['host_cards_per_s']),'dev',round(d['device_resident_cards_per_s']),d['oracle_equal'],d['oracle_checked'])
rr=json.loads([l for l in open('gpurun_out/${T}_bench_reference_n1.json') if l.startswith('{')][-1])
print('reference', rr['value'], rr['e2e']['value'], rr['cpu_baseline']['cores'], 'ratio', b['value']/rr['value'], 'e2e ratio', e['value']/rr['e2e']['value'])
PY
