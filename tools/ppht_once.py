#!/usr/bin/env python
"""One Canny + HoughLinesP(thr 0, 125/15) of the synthetic A4 card (profiling target)."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from omr_img_corrector_b200 import api, synth
g, a = synth.base_card(1)
ctx = api.Context(0)
bgr = ctx.rotate_mat(synth.to_bgr(g), -a, 1.0, api.INTER_LINEAR, api.WHITE, api.CLIP_DEFAULT)
e = ctx.canny(bgr, 50.0, 150.0)
t0 = time.perf_counter()
l = ctx.hough_lines_p(e, int(os.environ.get("THR", "0")), 125.0, 15.0)
print("edge px", int((e != 0).sum()), "lines", len(l), "seconds", time.perf_counter() - t0)
