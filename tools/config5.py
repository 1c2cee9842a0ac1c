#!/usr/bin/env python
"""BASELINE config 5 on N GPUs (run under torchrun): one huge pre-processed scan, +-MAX deg at
STEP deg, the angle range split over the ranks, per-angle std-devs all-gathered over NCCL and
the exact rule A applied on every rank.  Rank 0 also runs the whole sweep alone and asserts
the sharded result is bit-identical.  Prints one JSON line.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29511 tools/config5.py [--mul 4] [--max-angle 45] [--step 0.01]
"""
import argparse, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist
from omr_img_corrector_b200 import api, sharded, synth

ap = argparse.ArgumentParser()
ap.add_argument("--mul", type=int, default=4, help="A4@300dpi scaled by this factor (4 = 9921x14032 is config 5)")
ap.add_argument("--max-angle", type=int, default=45)
ap.add_argument("--step", type=float, default=0.01)
ap.add_argument("--check", type=int, default=1)
args = ap.parse_args()
rank, local, world = int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
ctx = api.Context(local)

# the A4 generator scaled xMUL with INTER_NEAREST: pre-processed on the GPU at A4, then replicated
g, a = synth.base_card(5)
bgr = ctx.rotate_mat(synth.to_bgr(g), -a, 1.0, api.INTER_LINEAR, api.WHITE, api.CLIP_DEFAULT)
b, s = ctx.preprocess(bgr, api.PRE_OMR, 0, 0)
big = np.repeat(np.repeat(b, args.mul, axis=0), args.mul, axis=1)
if args.mul == 4:
    big = np.ascontiguousarray(big[:14032, :9920])
    big = np.concatenate([big, big[:, -1:]], axis=1)          # 9921 x 14032
d_big = torch.from_numpy(big).to(dev)
h, w = big.shape


class DevImg:   # api._describe facade for a device-resident 8UC1 image
    pass


N, angles, M = api.sweep_matrices(w, h, args.max_angle, args.step, 1.0)


def sweep_fn(Ms):
    v, hh = ctx.projection_sweep_batch(d_big.data_ptr(), 1, w, h, Ms)
    return v[0], hh[0]


def run():
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    v, hh = sharded.sweep_sharded(sweep_fn, M, rank, world, None, dev)
    res = api.rule_a(hh, v, N, args.step)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    return time.perf_counter() - t0, v, hh, res


run()                       # warm-up: tables, planes, NCCL
t, v, hh, res = run()
tt = torch.tensor([t], dtype=torch.float64, device=dev)
if world > 1:
    dist.all_reduce(tt, op=dist.ReduceOp.MAX)
line = {"config": f"{w}x{h} scan, +-{args.max_angle} deg at {args.step} deg ({len(M)} angles), angle range split over {world} GPU(s), NCCL all-gather + rule A on every rank",
        "n_gpus": world, "seconds": float(tt.item()), "angle_evals_per_s": len(M) / float(tt.item()),
        "samples_per_s": len(M) * w * h / float(tt.item()), "angle": res.angle, "status": res.status, "skew": a}
if rank == 0 and args.check:
    t0 = time.perf_counter()
    v1, h1 = sweep_fn(M)
    line["single_gpu_seconds"] = time.perf_counter() - t0
    r1 = api.rule_a(h1, v1, N, args.step)
    line["identical_to_single_gpu"] = bool(np.array_equal(v1, v) and np.array_equal(h1, hh) and (r1.angle, r1.status) == (res.angle, res.status))
    assert line["identical_to_single_gpu"]
if rank == 0:
    print(json.dumps(line), flush=True)
if world > 1:
    dist.destroy_process_group()
