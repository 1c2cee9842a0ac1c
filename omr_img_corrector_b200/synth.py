"""Deterministic synthetic OMR answer cards.

numpy only.  The card is drawn un-skewed; the skew (rotate by -a about the
centre, INTER_LINEAR, white border, same size -- the input-synthesis recipe of
the reference's own tests, packages/lib/src/lib.rs:156-166) is applied by the
caller with whichever rotate implementation is under test.
"""
from __future__ import annotations

import numpy as np

A4_W, A4_H = 2480, 3508


def _ellipse_mask(ax, ay, half_w, half_h):
    ys, xs = np.mgrid[-half_h:half_h + 1, -half_w:half_w + 1]
    return (xs / ax) ** 2 + (ys / ay) ** 2 <= 1.0


_HW, _HH = 18, 12
_RING = _ellipse_mask(17.0, 11.0, _HW, _HH) & ~_ellipse_mask(15.0, 9.0, _HW, _HH)
_FILL = _ellipse_mask(14.0, 8.0, _HW, _HH)


def base_card(seed: int, w: int = A4_W, h: int = A4_H):
    """-> (gray uint8 [h, w], skew angle in degrees).

    Layout is defined on the A4@300dpi canvas (2480x3508) and scaled by
    k = w / 2480 for other sizes (k = 4 gives the A4@1200dpi scan of config 5).
    """
    rng = np.random.default_rng(seed)
    k = w / float(A4_W)
    W0, H0 = A4_W, A4_H
    img = np.full((H0, W0), 255, np.uint8)
    # 3-px black frame at 60 px inset
    img[60:63, 60:W0 - 60] = 0
    img[H0 - 63:H0 - 60, 60:W0 - 60] = 0
    img[60:H0 - 60, 60:63] = 0
    img[60:H0 - 60, W0 - 63:W0 - 60] = 0
    # 60 timing marks 60x22 at x=80, y=300+50r
    for r in range(60):
        img[300 + 50 * r:300 + 50 * r + 22, 80:140] = 0
    # 40 top marks 22x50 at y=120, x=300+50c
    for c in range(40):
        img[120:170, 300 + 50 * c:300 + 50 * c + 22] = 0
    # 60x40 bubble grid, row-major fill decisions
    fills = rng.random((60, 40)) < 0.35
    for r in range(60):
        cy = 311 + 50 * r
        for c in range(40):
            cx = 311 + 50 * c
            win = img[cy - _HH:cy + _HH + 1, cx - _HW:cx + _HW + 1]
            win[_RING] = 0
            if fills[r, c]:
                win[_FILL] = 0
    angle = round(float(rng.uniform(-14.0, 14.0)), 1)
    if (w, h) != (W0, H0):
        ki = int(round(k))
        if abs(k - ki) < 1e-9 and ki >= 1 and h == H0 * ki:
            img = np.repeat(np.repeat(img, ki, axis=0), ki, axis=1)   # INTER_NEAREST upscale
        else:
            ys = np.minimum((np.arange(h) * (H0 / float(h))).astype(np.int64), H0 - 1)
            xs = np.minimum((np.arange(w) * (W0 / float(w))).astype(np.int64), W0 - 1)
            img = img[ys][:, xs]
    return np.ascontiguousarray(img), angle


def to_bgr(gray: np.ndarray) -> np.ndarray:
    return np.ascontiguousarray(np.repeat(gray[:, :, None], 3, axis=2))


def skewed_card(seed: int, rotate_linear, w: int = A4_W, h: int = A4_H):
    """-> (BGR uint8 [h, w, 3], skew angle).  `rotate_linear(bgr, angle)` must be
    rotate_mat(bgr, angle, 1.0, INTER_LINEAR, white, DEFAULT)."""
    g, a = base_card(seed, w, h)
    return rotate_linear(to_bgr(g), -a), a
