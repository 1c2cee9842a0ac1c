#!/bin/bash
# 8 GPUs: full GPU suite (once), then N = 8, 4 bench lines as the driver launches them
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/r02z_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r02z_pytest_gpu.log
for N in 8 4; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2954$N bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/r02z_bench_n$N.json 2> gpurun_out/r02z_bench_n$N.err; echo "bench n$N rc=$?"
  python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/r02z_bench_n$N.json') if l.startswith('{')][-1])
e=d['e2e']
print('N=$N value', round(d['value'],1), 'ms/step', round(d['ms_per_step'],2), 'e2e files', round(e['value'],1), 'raw', round(e['raw_pixels']['value'],1), 'jpeg_out', round(e['jpeg_out']['value'],1), 'frac', round(d['roofline']['frac'],3))
print('  c5', {k:v for k,v in d['configs']['c5_huge_scan'].items() if k in ('seconds','single_gpu_seconds','identical_to_single_gpu','oracle_equal')})
PY
done
