#!/usr/bin/env python
"""config 2 with N A4 cards in flight (oics_result_from_edges_detection_batch): wall time per call; with
OICS_HOUGH_TRACE=1 the library prints its host phases."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from omr_img_corrector_b200 import api, synth
n = int(sys.argv[1]) if len(sys.argv) > 1 else 36
ctx = api.Context(0)
g, a = synth.base_card(0)
bgr = ctx.rotate_mat(synth.to_bgr(g), -a, 1.0, api.INTER_LINEAR, api.WHITE, api.CLIP_DEFAULT)
d = torch.from_numpy(bgr).cuda()[None].repeat(n, 1, 1, 1).contiguous()
for _ in range(2):
    ctx.result_from_edges_detection_batch(d, 125.0, 15.0)
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(3):
    r = ctx.result_from_edges_detection_batch(d, 125.0, 15.0)
dt = (time.perf_counter() - t0) / 3
print(f"{n} cards: {dt * 1e3:.1f} ms per call = {n / dt:.1f} cards/s")
