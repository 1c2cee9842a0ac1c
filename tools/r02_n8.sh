#!/bin/bash
set -u
mkdir -p gpurun_out
N=${1:-8}
timeout 300 python -m pytest tests -m gpu -q -x -k "two_gpus or two_devices" 2>&1 | tail -2
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/r02zz_bench_n$N.json 2> gpurun_out/r02zz_bench_n$N.err; echo "bench rc=$?"
python - <<PY
import json
b=json.loads([l for l in open('gpurun_out/r02zz_bench_n$N.json') if l.startswith('{')][-1])
e=b['e2e']; c=b['configs']['c5_huge_scan']
print('n_gpus', b['n_gpus'], 'value', round(b['value'],1), 'e2e', round(e['value'],1), 'raw', round(e['raw_pixels']['value'],1), 'jpeg_out', round(e['jpeg_out']['value'],1), 'c5', round(c['seconds'],4), c.get('identical_to_single_gpu'), c.get('oracle_equal'))
PY
tail -3 gpurun_out/r02zz_bench_n$N.err
