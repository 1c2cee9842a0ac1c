#!/bin/bash
# round 2, run 3 (2 GPUs): the full GPU test suite (incl. two devices in one process), then the bench as the
# driver launches it at N = 2 (config 5 sharded through oics_sweep_sharded + NCCL all-gather) and the reference arm
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/r02f_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r02f_pytest_gpu.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r02f_bench_n2.json 2> gpurun_out/r02f_bench_n2.err; echo "bench n2 rc=$?"; tail -3 gpurun_out/r02f_bench_n2.err
timeout 600 python bench.py --impl reference --gpus 1 --steps 5 --warmup 2 > gpurun_out/r02f_bench_reference.json 2> gpurun_out/r02f_bench_reference.err; echo "reference rc=$?"
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r02f_bench_n2.json') if l.startswith('{')][-1])
print('N=2 value', d['value'], 'e2e', d['e2e']['value'], 'c5', {k:v for k,v in d['configs']['c5_huge_scan'].items() if k!='what'})
r=json.loads([l for l in open('gpurun_out/r02f_bench_reference.json') if l.startswith('{')][-1])
print('reference value', r['value'], 'e2e', r['e2e']['value'], 'cores', r['cpu_baseline']['cores'], 'ms/step', r['ms_per_step'])
PY
