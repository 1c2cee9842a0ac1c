#!/usr/bin/env python
"""One A4-sized invocation of every secondary kernel family (device-resident inputs), with the library's own
per-kernel CUDA-event times and algorithmic bytes (oics_profile).  Run plain for the event times, or under
`ncu --set full -k regex:...` for the counters (tools/ncu_kernel_table.py merges the two)."""
import json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from omr_img_corrector_b200 import api, synth

out_path = sys.argv[1] if len(sys.argv) > 1 else None
ctx = api.Context(0)
g, a = synth.base_card(0)
bgr = ctx.rotate_mat(synth.to_bgr(g), -a, 1.0, api.INTER_LINEAR, api.WHITE, api.CLIP_DEFAULT)
H, W = bgr.shape[:2]
N = 8
d_cards = torch.from_numpy(bgr).cuda()[None].repeat(N, 1, 1, 1).contiguous()
side = 3310 * 4040 * 3
d_outs = [torch.empty(side, dtype=torch.uint8, device="cuda") for _ in range(N)]
d_gray = torch.from_numpy(np.ascontiguousarray(bgr[:, :, 0])).cuda()
file0 = np.frombuffer(ctx.jpeg_encode(bgr, 100), np.uint8)
pin = torch.empty((N, file0.size), dtype=torch.uint8).pin_memory()
files = []
for i in range(N):
    pin[i] = torch.from_numpy(file0.copy())
    files.append(pin[i].numpy())
jouts = [torch.empty(side // 2, dtype=torch.uint8).pin_memory().numpy() for _ in range(N)]


def run():
    ctx.correct_batch(d_cards, 15, 0.1, 0, 0, 125.0, 15.0, outs=d_outs)                   # P0, pack, sweep, finalize, W1 (NN)
    ctx.correct_batch_files(files, 15, 0.1, 0, 0, 125.0, 15.0, outs=jouts, jpeg_quality=100)   # both codecs
    ctx.rotate_mat(d_cards[0], -a, 1.0, api.INTER_LINEAR, api.WHITE, api.CLIP_DEFAULT)    # W1 (LINEAR), as the input synthesis
    e = ctx.canny(d_cards[0], 50.0, 150.0)                                                # H1 on BGR
    ctx.hough_peaks(ctx.hough_accumulate(e), 100)                                         # standard accumulator + peaks
    ctx.fft_image(d_gray)                                                                 # F1
    small = d_cards[:, :1150, :1240].contiguous()
    ctx.correct_batch(small, 45, 0.2, 248, 230, 150.0, 50.0, outs=d_outs)                 # the shrinking operating point


run()
ctx.profile(True)
run()
rep = ctx.profile_report()
ctx.profile(False)
txt = json.dumps(rep, indent=1)
if out_path:
    open(out_path, "w").write(txt)
print(txt)
