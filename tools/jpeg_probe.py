#!/usr/bin/env python
"""imwrite(JPEG, 100) of the deskewed A4 card: CUDA encoder (device-resident input, file bytes to a
pinned host buffer) beside the CPU oracle; parity asserted.  Prints one JSON object."""
import json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from omr_img_corrector_b200 import api, synth
from oracle import oracle as orc
g, a = synth.base_card(1)
ctx = api.Context(0)
bgr = ctx.rotate_mat(synth.to_bgr(g), -a, 1.0, api.INTER_LINEAR, api.WHITE, api.CLIP_DEFAULT)
card = ctx.rotate_mat(bgr, a, 1.0, api.INTER_NEAREST, api.WHITE, api.CLIP_CONTAIN)
d = torch.from_numpy(card).cuda()
out = torch.empty(64 << 20, dtype=torch.uint8).pin_memory().numpy()
got = ctx.jpeg_encode(d, 100, out=out)
best = 1e9
for _ in range(5):
    t0 = time.perf_counter(); ctx.jpeg_encode(d, 100, out=out); best = min(best, time.perf_counter() - t0)
t0 = time.perf_counter(); ref = orc.jpeg_encode(card, 100); tc = time.perf_counter() - t0
assert got == ref
print(json.dumps({"card": list(card.shape), "raw_bytes": card.size, "jpeg_bytes": len(got), "gpu_s": best, "cpu_oracle_s": tc,
                  "gpu_mpix_per_s": card.shape[0] * card.shape[1] / best / 1e6}))
