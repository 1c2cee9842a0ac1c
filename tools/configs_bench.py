#!/usr/bin/env python
"""BASELINE configs 1-3 on one synthetic A4 card: wall time of each estimator through the
C ABI (host buffers in, host results out) beside the CPU oracle port on the host cores.
Parity of every result is asserted.  Prints one JSON object."""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from omr_img_corrector_b200 import api, synth
from oracle import oracle as orc

g, a = synth.base_card(int(os.environ.get("CARD", "1")))
ctx = api.Context(0)
bgr = ctx.rotate_mat(synth.to_bgr(g), -a, 1.0, api.INTER_LINEAR, api.WHITE, api.CLIP_DEFAULT)
gray = orc.gray(bgr) if hasattr(orc, "gray") else None


def t(fn, n=3):
    fn()
    best = 1e9
    for _ in range(n):
        t0 = time.perf_counter()
        r = fn()
        best = min(best, time.perf_counter() - t0)
    return best, r


out = {"skew": a}
# config 1: projection method end to end (correct_default)
tg, rg = t(lambda: ctx.correct_default(bgr, 15, 0.1, 0, 0, 125.0, 15.0))
out["config1_correct_default"] = {"gpu_s": tg, "angle": rg[0]}
# the same with the caller's buffers pinned and reused (a fresh pageable 30 MB output array costs
# more in page faults than the whole GPU path)
import torch
pin_in = torch.from_numpy(bgr.copy()).pin_memory().numpy()
pin_out = torch.empty(int(np.hypot(*bgr.shape[:2]) + 3) ** 2 * 3, dtype=torch.uint8).pin_memory().numpy()
tg, rp = t(lambda: ctx.correct_default(pin_in, 15, 0.1, 0, 0, 125.0, 15.0, out=pin_out))
assert rp[0] == rg[0] and np.array_equal(rp[2], rg[2])
out["config1_correct_default"]["gpu_s_pinned_buffers"] = tg
tg, _ = t(lambda: ctx.correct_default(pin_in, 15, 0.1, 0, 0, 125.0, 15.0, want_image=False))
out["config1_correct_default"]["gpu_s_estimate_only"] = tg
# config 2: Hough method (Canny on BGR + HoughLinesP thr 0 + segment-angle mode)
tg, rg = t(lambda: ctx.result_from_edges_detection(bgr, 125.0, 15.0))
tc, rc = t(lambda: orc.get_result_from_edges_detection(bgr, 125.0, 15.0), n=1)
assert (rg.angle, rg.status) == (rc.angle, rc.status), (rg, rc)
out["config2_hough"] = {"gpu_s": tg, "cpu_s": tc, "angle": rg.angle, "status": rg.status}
tg, e = t(lambda: ctx.canny(bgr, 50.0, 150.0))
tc, ec = t(lambda: orc.canny(bgr, 50.0, 150.0), n=1)
assert np.array_equal(e, ec)
out["canny_bgr"] = {"gpu_s": tg, "cpu_s": tc, "edge_px": int((e != 0).sum())}
tg, l = t(lambda: ctx.hough_lines_p(e, 0, 125.0, 15.0))
tc, lc = t(lambda: orc.hough_lines_p(ec, 0, 125.0, 15.0), n=1)
assert np.array_equal(l, lc)
out["hough_lines_p_thr0"] = {"gpu_s": tg, "cpu_s": tc, "lines": int(len(l))}
tg, acc = t(lambda: ctx.hough_accumulate(e))
out["hough_accumulate"] = {"gpu_s": tg, "votes": int((e != 0).sum()) * 180}
# config 3: FFT method
g1 = ctx.rotate_mat(g, -a, 1.0, api.INTER_LINEAR, api.WHITE, api.CLIP_DEFAULT)
tg, rg = t(lambda: ctx.angle_with_fft(g1, 125.0, 150.0, 125.0, 15.0))
tc, rc = t(lambda: orc.get_angle_with_fft(g1, 125.0, 150.0, 125.0, 15.0), n=1)
out["config3_fft"] = {"gpu_s": tg, "cpu_s": tc, "angle": rg, "cpu_angle": rc}
tg, im = t(lambda: ctx.fft_image(g1))
out["fft_image"] = {"gpu_s": tg}
print(json.dumps(out, indent=1))
