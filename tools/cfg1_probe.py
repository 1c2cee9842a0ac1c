import time, numpy as np, torch, sys, os
sys.path.insert(0, os.getcwd())
from omr_img_corrector_b200 import api, synth
g, a = synth.base_card(1)
ctx = api.Context(0)
bgr = ctx.rotate_mat(synth.to_bgr(g), -a, 1.0, api.INTER_LINEAR, api.WHITE, api.CLIP_DEFAULT)
def t(fn, n=5):
    fn(); best = 1e9
    for _ in range(n):
        t0 = time.perf_counter(); fn(); best = min(best, time.perf_counter() - t0)
    return best * 1e3
print("pageable in, image out   ms", t(lambda: ctx.correct_default(bgr, 15, 0.1, 0, 0, 125.0, 15.0)))
print("pageable in, no image    ms", t(lambda: ctx.correct_default(bgr, 15, 0.1, 0, 0, 125.0, 15.0, want_image=False)))
pin = torch.from_numpy(bgr.copy()).pin_memory(); pb = pin.numpy()
print("pinned in, no image      ms", t(lambda: ctx.correct_default(pb, 15, 0.1, 0, 0, 125.0, 15.0, want_image=False)))
d = torch.from_numpy(bgr).cuda()
print("device in, no image      ms", t(lambda: ctx.correct_default(d, 15, 0.1, 0, 0, 125.0, 15.0, want_image=False)))
print("result_from_projection dev ms", t(lambda: ctx.result_from_projection(d, 15, 0.1, 0, 0)))
b, s = ctx.preprocess(bgr, api.PRE_OMR, 0, 0)
N, angles, M = api.sweep_matrices(2480, 3508, 15, 0.1, 1.0)
db = torch.from_numpy(b).cuda()
print("projection_sweep dev     ms", t(lambda: ctx.projection_sweep(db, M)))
print("preprocess dev->host     ms", t(lambda: ctx.preprocess(d, api.PRE_OMR, 0, 0)))
