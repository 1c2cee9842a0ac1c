#!/usr/bin/env python
"""Per-kernel device times (oics_profile, CUDA events) of one headline step: CARDS binary A4 cards resident in
HBM through oics_projection_sweep_batch + rule A.  Also the wall time of the call and of its parts."""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from omr_img_corrector_b200 import api, synth

cards_n = int(sys.argv[1]) if len(sys.argv) > 1 else 512
W, H = 2480, 3508
ctx = api.Context(0)
N, angles, M = api.sweep_matrices(W, H, 15, 0.1, 1.0)
g, a = synth.base_card(0)
bgr = ctx.rotate_mat(synth.to_bgr(g), -a, 1.0, api.INTER_LINEAR, api.WHITE, api.CLIP_DEFAULT)
b, s = ctx.preprocess(bgr, api.PRE_OMR, 0, 0)
cards = torch.from_numpy(b).cuda()[None].repeat(cards_n, 1, 1).contiguous()
out_v = torch.zeros((cards_n, 300), dtype=torch.float64).pin_memory().numpy()
out_h = torch.zeros((cards_n, 300), dtype=torch.float64).pin_memory().numpy()
for _ in range(3):
    ctx.projection_sweep_batch(cards.data_ptr(), cards_n, W, H, M, out=(out_v, out_h))
ctx.profile(True)
torch.cuda.synchronize()
t0 = time.perf_counter()
v, h = ctx.projection_sweep_batch(cards.data_ptr(), cards_n, W, H, M, out=(out_v, out_h))
t1 = time.perf_counter()
r = api.rule_a_batch(h, v, N, 0.1)
t2 = time.perf_counter()
rep = ctx.profile_report()
ctx.profile(False)
tot = sum(k["ms"] for k in rep)
print(json.dumps({"cards": cards_n, "call_ms": (t1 - t0) * 1e3, "rule_a_ms": (t2 - t1) * 1e3, "kernel_ms_total": tot,
                  "kernels": [{"kernel": k["kernel"], "launches": k["launches"], "ms": round(k["ms"], 3)} for k in rep]}, indent=1))
