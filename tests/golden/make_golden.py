#!/usr/bin/env python
"""Mints the golden vectors in tests/golden/ from the REAL OpenCV (Python cv2) driven
with the reference's exact call sequence, on cards from the reference's own dataset.

Run in the build container (needs cv2 and /root/reference); the output .npz files are
committed so that both the CPU oracle (`-m "not gpu"`) and the CUDA path (`-m gpu`) can be
checked where neither cv2 nor the reference tree exists (the GPU box).

    python tests/golden/make_golden.py

Every recorded quantity is produced by cv2 calls plus a literal Python transcription of
the Rust glue (count loops omr.rs:8-39, std-dev calculate.rs:2-23, argmax rules
omr.rs:147-221 / projection.rs:125-190, segment-angle rules omr.rs:257-301 /
hough.rs:48-89 / fft.rs:195-246) -- nothing from oracle/ or the product is imported.
"""
import ctypes
import hashlib
import math
import os
import sys

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
DATASET = "/root/reference/dataset/dataset"
WHITE = (255, 255, 255, 0)
_LIBM = ctypes.CDLL("libm.so.6")
_LIBM.atan2f.restype = ctypes.c_float
_LIBM.atan2f.argtypes = [ctypes.c_float, ctypes.c_float]


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def rot_matrix(w, h, angle, scale):
    # Point2f::new((w as f32)/2.0, (h as f32)/2.0)  (omr.rs:157-163)
    return cv2.getRotationMatrix2D((float(np.float32(w) / np.float32(2)), float(np.float32(h) / np.float32(2))),
                                   angle, scale)


def rotate_mat(img, angle, scale, flags, clip):
    """transfer::rotate_mat (transfer.rs:459-523)."""
    rows, cols = img.shape[:2]
    if clip == 0:
        M = rot_matrix(cols, rows, angle, scale)
        return cv2.warpAffine(img, M, (cols, rows), flags=flags, borderMode=cv2.BORDER_CONSTANT, borderValue=WHITE)
    r = angle * math.pi / 180.0
    rw = math.ceil(rows * abs(math.sin(r)) + cols * abs(math.cos(r)))
    rh = math.ceil(cols * abs(math.sin(r)) + rows * abs(math.cos(r)))
    M = cv2.getRotationMatrix2D((float(np.float32(math.ceil(rw / 2.0))), float(np.float32(math.ceil(rh / 2.0)))),
                                angle, scale)
    M[0, 2] += math.ceil((rw - cols) / 2.0)
    M[1, 2] += math.ceil((rh - rows) / 2.0)
    return cv2.warpAffine(img, M, (int(rw), int(rh)), flags=flags, borderMode=cv2.BORDER_CONSTANT,
                          borderValue=WHITE)


def std_dev(v):
    """calculate.rs:2-23: sequential f64 mean, sequential sum of (x-mean).powf(2), /n, powf(0.5)."""
    n = len(v)
    s = 0.0
    for x in v:
        s += float(x)
    mean = s / n
    acc = 0.0
    for x in v:
        d = float(x) - mean
        acc += d * d
    return math.sqrt(acc / n)


def preprocess_omr(bgr, max_w, max_h):
    """omr.rs:60-139."""
    h, w = bgr.shape[:2]
    ws = 1.0 if max_w <= 0 else max_w / float(w)
    hs = 1.0 if max_h <= 0 else max_h / float(h)
    s = ws if ws < hs else hs
    g = cv2.cvtColor(bgr, cv2.COLOR_RGB2GRAY)
    k = cv2.getStructuringElement(cv2.MORPH_ELLIPSE, (3, 3), (-1, -1))
    e = cv2.erode(g, k, anchor=(-1, -1), iterations=3, borderType=cv2.BORDER_CONSTANT)
    sc = cv2.resize(e, (int(w * s), int(h * s)), 0, 0, interpolation=cv2.INTER_AREA)
    _, b = cv2.threshold(sc, 127.0, 255.0, cv2.THRESH_BINARY)
    return b, s


def preprocess_projection(bgr, scale):
    """projection.rs:24-32 / transfer.rs:66-91."""
    h, w = bgr.shape[:2]
    if scale != 1.0:
        bgr = cv2.resize(bgr, (int(w * scale), int(h * scale)), 0, 0,
                         interpolation=cv2.INTER_AREA if scale < 1 else cv2.INTER_LINEAR)
    g = cv2.cvtColor(bgr, cv2.COLOR_RGB2GRAY)
    _, b = cv2.threshold(g, 127.0, 255.0, cv2.THRESH_BINARY)
    return b


def sweep(binary, max_angle, step, rot_scale):
    """omr.rs:140-208 / projection.rs:36-65: (N, rows[2N,h], cols[2N,w], v_sd, h_sd)."""
    h, w = binary.shape
    N = int(max_angle / step)
    rows, cols, v_sd, h_sd = [], [], [], []
    for deg in range(-N, N):
        M = rot_matrix(w, h, float(deg) * step, rot_scale)
        r = cv2.warpAffine(binary, M, (w, h), flags=0, borderMode=cv2.BORDER_CONSTANT, borderValue=WHITE)
        z = r == 0
        rc, cc = z.sum(axis=1).astype(np.int32), z.sum(axis=0).astype(np.int32)
        rows.append(rc)
        cols.append(cc)
        h_sd.append(std_dev(rc))
        v_sd.append(std_dev(cc))
    return N, np.stack(rows), np.stack(cols), np.array(v_sd), np.array(h_sd)


def rule_a(h, v, N, step):
    mh = mv = 0.0
    ph = pv = 1
    most = []
    for i, deg in enumerate(range(-N, N)):
        a = float(deg) * step
        if mh < h[i]:
            mh, mv, ph, pv, most = h[i], v[i], 1, 1, [a]
        elif mh == h[i]:
            ph += 1
            if mv < v[i]:
                pv, mv, most = 1, v[i], [a]
            elif mv == v[i]:
                pv += 1
                most.append(a)
    if ph == 1 and pv == 1:
        return most[0], 0, most
    if len(most) == 1:
        return most[0], 1, most
    return 0.0, 2, most


def rule_b(v, h, N, step):
    n = 2 * N
    vmax, hmax, vl, hl = v[0], h[0], [0], [0]
    for i in range(n):
        if v[i] > vmax:
            vmax, vl = v[i], [i]
        elif v[i] == vmax:
            vl.append(i)
        if h[i] > hmax:
            hmax, hl = h[i], [i]
        elif h[i] == hmax:
            hl.append(i)
    if len(vl) == 1 and len(hl) == 1 and vl[0] == hl[0]:
        idx = vl[0]
    else:
        sdp, idx = 0.0, -1
        for c in sorted(set(vl + hl)):
            cur = v[c] * v[c] + h[c] * h[c]
            if sdp < cur:
                sdp, idx = cur, c
        if idx < 0:
            idx = n // 2
    return (float(idx) - float(N)) * step


def seg_angle_f32(l):
    x1, y1, x2, y2 = [np.float32(t) for t in l]
    # f32::atan2 = the platform libm's atan2f (glibc's here, correctly rounded) -- numpy's SIMD
    # float32 arctan2 is 1 ulp off on ~1/3 of the segments and is not used.
    t = np.float32(_LIBM.atan2f(float(np.float32(y2 - y1)), float(np.float32(x2 - x1))))
    a = np.float32(np.float32(t * np.float32(180.0)) / np.float32(math.pi))
    return np.float32(math.fmod(float(np.float32(a)), 45.0))


def lines_angle_omr(lines):
    ang = [float(seg_angle_f32(l)) for l in lines]
    target, best, cand = ang[0], 0, []
    for i in range(len(ang)):
        c = sum(1 for j in range(len(ang)) if abs(ang[i] - ang[j]) < 0.1)
        if c > best:
            best, target, cand = c, ang[i], [ang[i]]
        elif c == best:
            cand.append(ang[i])
    status = 2 if not cand else 0 if len(cand) == 1 else 1
    return target, status, cand


def lines_angle_hough(lines):
    ang = [seg_angle_f32(l) for l in lines]
    target, best = ang[0], 0
    for i in range(len(ang)):
        c = sum(1 for j in range(len(ang)) if abs(np.float32(ang[i] - ang[j])) < np.float32(0.1))
        if c > best:
            best, target = c, ang[i]
    return float(target)


def lines_angle_fft(lines):
    best_angle, max_votes, n = 0.0, 0, len(lines)
    for i in range(n):
        l = lines[i]
        raw = math.atan2(float(l[3]) - float(l[1]), float(l[2]) - float(l[0])) * 180.0 / math.pi
        folded = raw + 90.0 if raw < -45.0 else raw - 90.0 if raw > 45.0 else raw
        votes = sum(1 for j in range(n) if j != i and abs(raw - folded) < 0.1)     # fft.rs:231 index bug
        if votes > max_votes:
            max_votes, best_angle = votes, folded
    return best_angle


def hough_p(edges, thr, min_len, max_gap):
    l = cv2.HoughLinesP(edges, 1.0, math.pi / 180.0, thr, minLineLength=min_len, maxLineGap=max_gap)
    return np.zeros((0, 4), np.int32) if l is None else l.reshape(-1, 4).astype(np.int32)


def fft_image(gray):
    """fft.rs:124-141 with cv2 (float32 cv::dft)."""
    src = gray.astype(np.float64)
    f32 = (src * (1.0 / 255.0)).astype(np.float32)                 # convertTo(CV_32F, 1/255)
    planes = cv2.merge([f32, np.zeros_like(f32)])
    F = cv2.dft(planes, flags=cv2.DFT_COMPLEX_OUTPUT | cv2.DFT_COMPLEX_INPUT | cv2.DFT_SCALE)
    re, im = cv2.split(F)
    h, w = gray.shape

    def shift(p):
        q = p.copy()
        cx, cy = w // 2, h // 2
        q[0:cy, 0:cx] = p[cy:2 * cy, cx:2 * cx]
        q[cy:2 * cy, cx:2 * cx] = p[0:cy, 0:cx]
        q[0:cy, cx:2 * cx] = p[cy:2 * cy, 0:cx]
        q[cy:2 * cy, 0:cx] = p[0:cy, cx:2 * cx]
        return q

    mag = cv2.magnitude(shift(re), shift(im))

    def correction(a):
        mn, mx, _, _ = cv2.minMaxLoc(a)
        b = (a.astype(np.float64) - mn).astype(np.float32)
        return (b.astype(np.float64) * (1.0 / (mx - mn))).astype(np.float32)

    m = (correction(mag).astype(np.float64) * 255.0).astype(np.float32)
    lg = (m.astype(np.float64) + 1.0 / 255.0).astype(np.float32)
    lg = cv2.log(lg)
    out = correction(lg)
    return np.clip(np.rint(out.astype(np.float64) * 255.0), 0, 255).astype(np.uint8)


def make_card(name, skew):
    """Dataset card halved (INTER_AREA) to keep the fixture small, skewed with the reference's
    input-synthesis recipe (lib.rs:156-166: rotate by -a, INTER_LINEAR, DEFAULT clip, white)."""
    img = cv2.imread(os.path.join(DATASET, name), cv2.IMREAD_COLOR)
    h, w = img.shape[:2]
    img = cv2.resize(img, (w // 2, h // 2), interpolation=cv2.INTER_AREA)
    g = cv2.cvtColor(img, cv2.COLOR_BGR2GRAY)
    bgr = cv2.cvtColor(g, cv2.COLOR_GRAY2RGB)                       # lib.rs:168-176: gray -> 3 equal channels
    return rotate_mat(bgr, -skew, 1.0, 1, 0)


def golden_for(name, skew, out):
    bgr = make_card(name, skew)
    gray = cv2.cvtColor(bgr, cv2.COLOR_RGB2GRAY)
    d = {"gray_in": bgr[:, :, 0].copy(), "skew": np.float64(skew), "cv2_version": np.array(cv2.__version__),
         "ipp": np.array(bool(cv2.ipp.useIPP()))}
    # P0 + sweep at the reference's default operating point (lib.rs:92-97) scaled to the halved card
    for tag, (max_angle, step, mw, mh) in {"def": (45, 0.5, 124, 115), "full": (10, 1.0, 0, 0)}.items():
        b, s = preprocess_omr(bgr, mw, mh)
        N, rows, cols, v, h = sweep(b, max_angle, step, s)
        a, st, cand = rule_a(h, v, N, step)
        d[f"{tag}_params"] = np.array([max_angle, step, mw, mh], np.float64)
        d[f"{tag}_bin"] = np.packbits(b == 0)
        d[f"{tag}_bin_shape"] = np.array(b.shape)
        d[f"{tag}_scale"] = np.float64(s)
        d[f"{tag}_rows"], d[f"{tag}_cols"] = rows[::7], cols[::7]
        d[f"{tag}_v_sd"], d[f"{tag}_h_sd"] = v, h
        d[f"{tag}_rule_a"] = np.array([a, st], np.float64)
        d[f"{tag}_rule_a_cand"] = np.array(cand, np.float64)
    # P0' + rule B (projection.rs path), resize_scale 0.4
    b = preprocess_projection(bgr, 0.4)
    N, rows, cols, v, h = sweep(b, 45, 0.5, 1.0)
    d["proj_bin"] = np.packbits(b == 0)
    d["proj_bin_shape"] = np.array(b.shape)
    d["proj_v_sd"], d["proj_h_sd"] = v, h
    d["proj_rule_b"] = np.float64(rule_b(v, h, N, 0.5))
    # H1 / H2 / H3
    e_gray = cv2.Canny(gray, 50.0, 150.0, apertureSize=3, L2gradient=False)
    e_bgr = cv2.Canny(bgr, 50.0, 150.0, apertureSize=3, L2gradient=False)
    d["canny_gray"], d["canny_bgr"] = np.packbits(e_gray != 0), np.packbits(e_bgr != 0)
    for thr, ml, mg in [(0, 60.0, 8.0), (0, 75.0, 25.0), (100, 60.0, 8.0)]:
        L = hough_p(e_gray, thr, ml, mg)
        key = f"lines_{thr}_{int(ml)}_{int(mg)}"
        d[key] = L
        if len(L):
            a, st, cand = lines_angle_omr(L)
            d[key + "_omr"] = np.array([a, st], np.float64)
            d[key + "_omr_cand"] = np.array(cand, np.float64)
            d[key + "_hough"] = np.float64(lines_angle_hough(L))
            d[key + "_fft"] = np.float64(lines_angle_fft(L))
    # F1 (cv::dft float32; compare with +-1 LSB tolerance) and the FFT estimator's pieces
    log8 = fft_image(gray)
    d["fft_log8"] = log8
    e_fft = cv2.Canny(log8, 125.0, 150.0, apertureSize=3, L2gradient=False)
    d["fft_canny"] = np.packbits(e_fft != 0)
    d["fft_lines"] = hough_p(e_fft, 100, 60.0, 8.0)
    # W1: hashes of the four final-correction variants at the recovered angle
    ang = float(d["def_rule_a"][0]) if d["def_rule_a"][1] == 0 else skew
    d["warp_angle"] = np.float64(ang)
    for flags in (0, 1):
        for clip in (0, 1):
            r = rotate_mat(bgr, ang, 1.0, flags, clip)
            d[f"warp_{flags}_{clip}_shape"] = np.array(r.shape)
            d[f"warp_{flags}_{clip}_sha"] = np.array(sha(r))
    np.savez_compressed(out, **d)
    print(out, os.path.getsize(out) // 1024, "KiB", "angle", d["def_rule_a"], "skew", skew)


if __name__ == "__main__":
    if not os.path.isdir(DATASET):
        sys.exit("needs /root/reference/dataset/dataset")
    golden_for("image001.jpg", 3.7, os.path.join(HERE, "card_image001.npz"))
    golden_for("SCN00025_2.jpg", -11.3, os.path.join(HERE, "card_scn00025.npz"))
