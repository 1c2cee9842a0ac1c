#!/bin/bash
# Multi-GPU checks: bench at N ranks and config 5 (sharded angle range over NCCL).
set -u
N=${1:-2}
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29501 bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_n$N.err; cat gpurun_out/bench_n$N.json
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29502 tools/config5.py > gpurun_out/config5_n$N.json 2> gpurun_out/config5_n$N.err; echo "config5 rc=$?"; tail -5 gpurun_out/config5_n$N.err; cat gpurun_out/config5_n$N.json
