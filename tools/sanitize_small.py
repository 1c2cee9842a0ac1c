#!/usr/bin/env python
"""Small invocations of every kernel family, for compute-sanitizer (memcheck / racecheck / synccheck):

    compute-sanitizer --tool racecheck python tools/sanitize_small.py

Sizes are chosen so the word sweep kernel (sweepw_kernel: double-buffered row tables and row sums), the
general and the small-card sweep kernels, Canny (cooperative grid barriers), the PPHT cluster kernel (DSMEM +
barrier.cluster ordering, several clusters in flight), the FFT, the warp and both JPEG codecs all run, each
checked against the CPU restatement so that a sanitizer-clean run is also a correct one."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from omr_img_corrector_b200 import api, synth  # noqa: E402
from oracle import oracle as orc  # noqa: E402  (tools/ is test infrastructure)

only = set(sys.argv[1:])


def want(name):
    return not only or name in only


ctx = api.Context(0)
g, a = synth.base_card(3)
crop = np.ascontiguousarray(g[250:850, 250:810])                      # 600 x 560
bgr = orc.rotate_mat(synth.to_bgr(crop), -2.5, 1.0, 1, orc.WHITE, 0)
if want("sweep"):
    b = orc.preprocess_projection(bgr, 1.0)
    N, ang, M = api.sweep_matrices(b.shape[1], b.shape[0], 3, 0.5, 1.0)     # word kernel
    Mg = np.stack([api.rotation_matrix_2d(280.0, 300.0, x, 0.7) for x in (-30.0, 33.0)])   # general kernel
    Ms = np.concatenate([M, Mg])
    v, h = ctx.projection_sweep(b, Ms)
    rv, rh = orc.projection_sweep(b, Ms, threads=4)
    assert np.array_equal(v, rv) and np.array_equal(h, rh)
    bs, s = orc.preprocess_omr(bgr, 124, 115)                                 # small-card kernel, scaled rotation
    N, ang, M = api.sweep_matrices(bs.shape[1], bs.shape[0], 10, 1.0, s)
    v, h = ctx.projection_sweep(bs, M)
    rv, rh = orc.projection_sweep(bs, M, threads=4)
    assert np.array_equal(v, rv) and np.array_equal(h, rh)
    print("sweep ok", flush=True)
if want("canny") or want("hough"):
    e = ctx.canny(bgr, 50.0, 150.0)
    assert np.array_equal(e, orc.canny(bgr, 50.0, 150.0))
    print("canny ok", flush=True)
if want("hough"):
    small = np.ascontiguousarray(bgr[:300, :260])
    cards = np.stack([small, np.ascontiguousarray(small[::-1]), np.ascontiguousarray(small[:, ::-1])])
    ang, rs, st = ctx.result_from_edges_detection_batch(cards, 50.0, 5.0)     # three clusters in flight
    for i in range(3):
        o = orc.get_result_from_edges_detection(cards[i], 50.0, 5.0)
        assert st[i] == 0 and (o.angle, o.status) == (ang[i], rs[i])
    acc = ctx.hough_accumulate(e)
    assert np.array_equal(ctx.hough_peaks(acc, 60), orc.hough_peaks(acc, 60))
    print("hough ok", flush=True)
if want("fft"):
    gray = orc.gray(bgr)[:200, :232]
    gray = np.ascontiguousarray(gray)
    d = np.abs(ctx.fft_image(gray)[1].astype(np.int16) - orc.fft_image(gray)[1].astype(np.int16))
    assert d.max() <= 1
    print("fft ok", flush=True)
if want("warp"):
    for flags in (0, 1):
        assert np.array_equal(ctx.rotate_mat(bgr, 3.3, 1.0, flags, api.WHITE, 1), orc.rotate_mat(bgr, 3.3, 1.0, flags, orc.WHITE, 1))
    assert np.array_equal(ctx.preprocess(bgr, api.PRE_OMR, 0, 0)[0], orc.preprocess_omr(bgr, 0, 0)[0])
    assert np.array_equal(ctx.preprocess(bgr, api.PRE_OMR, 124, 115)[0], orc.preprocess_omr(bgr, 124, 115)[0])
    print("warp + pre-processing ok", flush=True)
if want("jpeg"):
    small = np.ascontiguousarray(bgr[:232, :312])
    data = ctx.jpeg_encode(small, 100)
    assert data == orc.jpeg_encode(small, 100)
    assert np.array_equal(ctx.jpeg_decode(data), orc.jpeg_decode(data))
    outs = [np.zeros(1 << 20, np.uint8) for _ in range(3)]
    files = [np.frombuffer(data, np.uint8)] * 3
    r = ctx.correct_batch_files(files, 5, 0.5, 0, 0, 50.0, 5.0, outs=outs, jpeg_quality=100)
    assert not r[2].any()
    print("jpeg ok", flush=True)
print("sanitize_small: all ok")
