#!/usr/bin/env python
"""One imread + one imwrite of an A4 card alone on the device (warm), with the per-kernel event times of
oics_profile: the codec kernels of the file-to-file path (jpegdec.cu, jpeg.cu) in isolation."""
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
fixed = ctx.rotate_mat(bgr, a, 1.0, api.INTER_NEAREST, api.WHITE, api.CLIP_CONTAIN)     # what imwrite gets
d_fixed = torch.from_numpy(fixed).cuda()
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 5
for _ in range(2):
    ctx.jpeg_decode(data, out=out)
    ctx.jpeg_encode(d_fixed, 100)
ctx.profile(True)
for _ in range(reps):
    ctx.jpeg_decode(data, out=out)
    ctx.jpeg_encode(d_fixed, 100)
rep = ctx.profile_report()
ctx.profile(False)
print(json.dumps({"file_bytes": int(data.size), "out_shape": list(fixed.shape),
                  "us_per_card": {k["kernel"]: round(1000 * k["ms"] / reps, 1) for k in rep}}))
