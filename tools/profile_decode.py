#!/usr/bin/env python
"""Per-kernel device times of the JPEG decoder on one A4 card's file, alone on the device (oics_profile)."""
import json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from omr_img_corrector_b200 import api, synth
ctx = api.Context(0)
g, a = synth.base_card(0)
bgr = ctx.rotate_mat(synth.to_bgr(g), -a, 1.0, api.INTER_LINEAR, api.WHITE, api.CLIP_DEFAULT)
data = np.frombuffer(ctx.jpeg_encode(bgr, 100), np.uint8)
out = torch.empty(bgr.size, dtype=torch.uint8, device="cuda")
for _ in range(2):
    ctx.jpeg_decode(data, out=out)
ctx.profile(True)
ctx.jpeg_decode(data, out=out)
rep = ctx.profile_report()
ctx.profile(False)
print(json.dumps({"file_bytes": int(data.size), "kernels": [{"kernel": k["kernel"], "launches": k["launches"], "us": round(1000 * k["ms"], 1)} for k in rep]}, indent=1))
