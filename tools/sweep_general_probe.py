#!/usr/bin/env python
"""The general sweep kernel on BASELINE config 5's scan at large angles (the part of the +-45 degree sweep the word
kernel cannot take): device time per angle, and a target for ncu."""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from omr_img_corrector_b200 import api, synth
ctx = api.Context(0)
g, a = synth.base_card(5)
bgr = ctx.rotate_mat(synth.to_bgr(g), -a, 1.0, api.INTER_LINEAR, api.WHITE, api.CLIP_DEFAULT)
b, s = ctx.preprocess(bgr, api.PRE_OMR, 0, 0)
big = np.repeat(np.repeat(b, 4, axis=0), 4, axis=1)
big = np.ascontiguousarray(big[:14032, :9920])
big = np.ascontiguousarray(np.concatenate([big, big[:, -1:]], axis=1))
d = torch.from_numpy(big).cuda()
h, w = big.shape
angles = [float(x) for x in (sys.argv[1:] or [20.0, 25.0, 30.0, 35.0, 40.0, 44.99, -22.5, -44.0])]
M = np.stack([api.rotation_matrix_2d(np.float32(w) / np.float32(2), np.float32(h) / np.float32(2), x, 1.0) for x in angles])
ctx.projection_sweep_batch(d.data_ptr(), 1, w, h, M)
ctx.profile(True)
torch.cuda.synchronize()
t0 = time.perf_counter()
ctx.projection_sweep_batch(d.data_ptr(), 1, w, h, M)
t1 = time.perf_counter()
rep = ctx.profile_report()
ctx.profile(False)
print(json.dumps({"angles": angles, "call_ms": (t1 - t0) * 1e3, "samples": len(angles) * w * h,
                  "kernels": [{"kernel": k["kernel"], "launches": k["launches"], "ms": round(k["ms"], 3)} for k in rep]}))
