#!/usr/bin/env python
"""imread of an A4 card's JPEG (quality 100, written by the CUDA encoder): CUDA decoder (file bytes in
pinned host memory, pixels left on the device) beside the CPU oracle and cv2; parity asserted."""
import json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from omr_img_corrector_b200 import api, synth
from oracle import oracle as orc
g, a = synth.base_card(1)
ctx = api.Context(0)
bgr = ctx.rotate_mat(synth.to_bgr(g), -a, 1.0, api.INTER_LINEAR, api.WHITE, api.CLIP_DEFAULT)
data = ctx.jpeg_encode(bgr, 100)
pin = torch.empty(len(data), dtype=torch.uint8).pin_memory()
pin.numpy()[:] = np.frombuffer(data, np.uint8)
dev = torch.empty(bgr.size, dtype=torch.uint8, device="cuda")
got = ctx.jpeg_decode(pin.numpy(), out=dev)
best = 1e9
for _ in range(5):
    t0 = time.perf_counter(); ctx.jpeg_decode(pin.numpy(), out=dev); best = min(best, time.perf_counter() - t0)
t0 = time.perf_counter(); ref = orc.jpeg_decode(data); tc = time.perf_counter() - t0
assert np.array_equal(got.cpu().numpy(), ref)
out = {"card": list(bgr.shape), "jpeg_bytes": len(data), "raw_bytes": bgr.size, "gpu_s": best, "cpu_oracle_s": tc,
       "gpu_mpix_per_s": bgr.shape[0] * bgr.shape[1] / best / 1e6}
try:
    import cv2
    t0 = time.perf_counter(); c = cv2.imdecode(np.frombuffer(data, np.uint8), cv2.IMREAD_COLOR); out["cv2_s"] = time.perf_counter() - t0
    out["equals_cv2"] = bool(np.array_equal(c, ref))
except ImportError:
    pass
print(json.dumps(out))
