"""CPU oracle for the deskew hot path -- TEST INFRASTRUCTURE ONLY.

Python front end over ``oracle/_ref/liboics_oracle.so`` (the C restatement in
``oics_oracle.c`` / ``oics_oracle_lines.c``) plus a numpy restatement of the
FFT image.  Only ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` may import this
module; the product package never does.

Every function cites the reference file:line it follows (paths relative to the
reference repo root).  How the oracle is pinned is described in the header of
``oics_oracle.c``.
"""
from __future__ import annotations

import ctypes as C
import math
import os
import subprocess
from dataclasses import dataclass, field

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_ref", "liboics_oracle.so")

BELIEVED, NEED_CHECK, NOT_A_RESULT = 0, 1, 2
CLIP_DEFAULT, CLIP_CONTAIN = 0, 1
INTER_NEAREST, INTER_LINEAR = 0, 1
WHITE = (255, 255, 255, 0)


def build(force: bool = False) -> str:
    """Compile the C restatement into oracle/_ref/ (recipe: oracle/Makefile)."""
    srcs = [os.path.join(_HERE, f) for f in ("oics_oracle.c", "oics_oracle_lines.c", "oics_oracle_jpeg.c", "oics_oracle_jpegdec.c", "Makefile")]
    if (not force and os.path.exists(_SO)
            and all(os.path.getmtime(_SO) >= os.path.getmtime(s) for s in srcs)):
        return _SO
    env = dict(os.environ)
    env["CC"] = "/usr/bin/gcc" if os.path.exists("/usr/bin/gcc") else "gcc"
    subprocess.check_call(["make", "-C", _HERE, "-B"], env=env, stdout=subprocess.DEVNULL)
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_SO)
        _lib.orc_std_dev_i32.restype = C.c_double
        _lib.orc_projection_scale.restype = C.c_double
        _lib.orc_lines_angle_fft.restype = C.c_double
    return _lib


def _p(a, t=C.c_uint8):
    return a.ctypes.data_as(C.POINTER(t))


def _img(a):
    a = np.ascontiguousarray(a, dtype=np.uint8)
    h, w = a.shape[:2]
    ch = 1 if a.ndim == 2 else a.shape[2]
    return a, w, h, ch


# --------------------------------------------------------------------------- geometry
def rotation_matrix_2d(cx, cy, angle, scale):
    """cv::getRotationMatrix2D (omr.rs:157-163); centre rounded to f32 first."""
    M = np.zeros(6, np.float64)
    lib().orc_rotation_matrix_2d(C.c_float(cx), C.c_float(cy), C.c_double(angle), C.c_double(scale),
                                 _p(M, C.c_double))
    return M


def rotate_geometry(cols, rows, angle, scale, clip):
    """transfer::rotate_mat geometry (transfer.rs:459-523) -> (M[6], dw, dh)."""
    M = np.zeros(6, np.float64)
    dw, dh = C.c_int(), C.c_int()
    lib().orc_rotate_geometry(cols, rows, C.c_double(angle), C.c_double(scale), clip,
                              _p(M, C.c_double), C.byref(dw), C.byref(dh))
    return M, dw.value, dh.value


def sweep_angles(max_angle, step):
    """Half-open grid -N..N, angle = deg as f64 * step (omr.rs:140-161)."""
    N = lib().orc_angle_half_range(C.c_uint(max_angle), C.c_double(step))
    return N, [float(d) * step for d in range(-N, N)]


def sweep_matrices(w, h, max_angle, step, scale):
    N, angles = sweep_angles(max_angle, step)
    cx, cy = np.float32(w) / np.float32(2.0), np.float32(h) / np.float32(2.0)
    Ms = np.stack([rotation_matrix_2d(cx, cy, a, scale) for a in angles]) if angles else np.zeros((0, 6))
    return N, angles, np.ascontiguousarray(Ms)


# --------------------------------------------------------------------------- warps
def warp_affine(src, M, dsize, flags=INTER_NEAREST, border=WHITE):
    """cv::warpAffine(flags 0|1, BORDER_CONSTANT) (omr.rs:165-173, transfer.rs:477-485)."""
    src, w, h, ch = _img(src)
    dw, dh = dsize
    dst = np.empty((dh, dw) if src.ndim == 2 else (dh, dw, ch), np.uint8)
    M = np.ascontiguousarray(M, np.float64).reshape(6)
    b = np.array([int(v) & 255 for v in border], np.uint8)
    fn = lib().orc_warp_affine_nn if flags == INTER_NEAREST else lib().orc_warp_affine_linear
    fn(_p(src), w, h, w * ch, ch, _p(dst), dw, dh, dw * ch, _p(M, C.c_double), _p(b))
    return dst


def rotate_mat(src, angle, scale=1.0, flags=INTER_NEAREST, border=WHITE, clip=CLIP_DEFAULT):
    """transfer::rotate_mat (transfer.rs:459-523)."""
    src, w, h, ch = _img(src)
    M, dw, dh = rotate_geometry(w, h, angle, scale, clip)
    return warp_affine(src, M, (dw, dh), flags, border)


# --------------------------------------------------------------------------- pre-processing
def gray(bgr):
    """cvtColor(COLOR_RGB2GRAY) on BGR bytes (omr.rs:90, transfer.rs:283-290)."""
    bgr, w, h, ch = _img(bgr)
    assert ch == 3
    out = np.empty((h, w), np.uint8)
    lib().orc_gray(_p(bgr), w, h, w * 3, _p(out), w)
    return out


def erode3(g, iters=3):
    """erode, 3x3 MORPH_ELLIPSE, 3 iterations (omr.rs:96-112)."""
    g, w, h, ch = _img(g)
    out = np.empty_like(g)
    lib().orc_erode_cross(_p(g), w, h, w, _p(out), w, iters)
    return out


def resize_area(img, dw, dh):
    """cv::resize(INTER_AREA), shrinking (omr.rs:116-126, transfer.rs:72-87)."""
    img, w, h, ch = _img(img)
    out = np.empty((dh, dw) if img.ndim == 2 else (dh, dw, ch), np.uint8)
    rc = lib().orc_resize_area(_p(img), w, h, w * ch, ch, _p(out), dw, dh, dw * ch)
    if rc != 0:
        raise ValueError("enlarging resize is not restated")
    return out


def threshold(g):
    """threshold(127, 255, THRESH_BINARY) (omr.rs:131-137)."""
    g, w, h, ch = _img(g)
    out = np.empty_like(g)
    lib().orc_threshold(_p(g), w, h, w, _p(out), w)
    return out


def projection_scale(w, h, max_w, max_h):
    """omr.rs:60-82."""
    return lib().orc_projection_scale(w, h, max_w, max_h)


def preprocess_omr(bgr, max_w, max_h):
    """P0: omr.rs:60-139 -> (binary, scale)."""
    bgr, w, h, ch = _img(bgr)
    s = projection_scale(w, h, max_w, max_h)
    e = erode3(gray(bgr), 3)
    sc = resize_area(e, int(float(w) * s), int(float(h) * s))
    return threshold(sc), s


def preprocess_projection(bgr, resize_scale):
    """P0': projection.rs:24-32 (scale colour image, gray, threshold; no erode)."""
    bgr, w, h, ch = _img(bgr)
    if resize_scale != 1.0:
        if resize_scale > 1.0:
            raise ValueError("enlarging resize is not restated")
        bgr = resize_area(bgr, int(float(w) * resize_scale), int(float(h) * resize_scale))
    return threshold(gray(bgr))


# --------------------------------------------------------------------------- counts / std-dev
def projection_counts(img):
    """omr::get_mat_projection_data (omr.rs:8-39) -> (row_counts, col_counts)."""
    img, w, h, ch = _img(img)
    row = np.empty(h, np.int32)
    col = np.empty(w, np.int32)
    lib().orc_projection_counts(_p(img), w, h, w, _p(row, C.c_int32), _p(col, C.c_int32))
    return row, col


def std_dev(v):
    """calculate::get_standard_deviation (calculate.rs:13-23)."""
    v = np.ascontiguousarray(v, np.int32)
    return lib().orc_std_dev_i32(_p(v, C.c_int32), len(v))


def projection_std_devs(img):
    """transfer::get_projection_standard_deviations (transfer.rs:527-536) -> (v_sd, h_sd)."""
    row, col = projection_counts(img)
    return std_dev(col), std_dev(row)


def projection_sweep(binary, Ms, want_counts=False, threads=1):
    """The hot loop (omr.rs:153-208 / projection.rs:47-65) for forward matrices Ms."""
    binary, w, h, ch = _img(binary)
    Ms = np.ascontiguousarray(Ms, np.float64).reshape(-1, 6)
    n = len(Ms)
    v = np.zeros(n, np.float64)
    hh = np.zeros(n, np.float64)
    rows = np.zeros((n, h), np.int32) if want_counts else None
    cols = np.zeros((n, w), np.int32) if want_counts else None
    lib().orc_projection_sweep(_p(binary), w, h, w, _p(Ms, C.c_double), n, _p(v, C.c_double),
                               _p(hh, C.c_double),
                               _p(rows, C.c_int32) if want_counts else None,
                               _p(cols, C.c_int32) if want_counts else None, threads)
    return (v, hh, rows, cols) if want_counts else (v, hh)


@dataclass
class OmrResult:
    """omr::OmrResult (omr.rs:41-50)."""
    angle: float
    status: int
    candidates: list = field(default_factory=list)


def rule_a(h_sd, v_sd, N, step):
    """omr.rs:147-221."""
    h_sd = np.ascontiguousarray(h_sd, np.float64)
    v_sd = np.ascontiguousarray(v_sd, np.float64)
    cap = max(1, 2 * N)
    cand = np.zeros(cap, np.float64)
    angle, status, nc = C.c_double(), C.c_int(), C.c_int()
    lib().orc_rule_a(_p(h_sd, C.c_double), _p(v_sd, C.c_double), N, C.c_double(step),
                     C.byref(angle), C.byref(status), _p(cand, C.c_double), cap, C.byref(nc))
    return OmrResult(angle.value, status.value, list(cand[:nc.value]))


def rule_b(v_sd, h_sd, N, step):
    """projection.rs:125-190 -> angle."""
    v_sd = np.ascontiguousarray(v_sd, np.float64)
    h_sd = np.ascontiguousarray(h_sd, np.float64)
    idx = lib().orc_rule_b_index(_p(v_sd, C.c_double), _p(h_sd, C.c_double), len(v_sd))
    return (float(idx) - float(N)) * step


def get_result_from_projection(bgr, max_angle, step, max_w, max_h, threads=1):
    """omr::get_result_from_projection (omr.rs:52-227)."""
    binary, s = preprocess_omr(bgr, max_w, max_h)
    h, w = binary.shape
    N, angles, Ms = sweep_matrices(w, h, max_angle, step, s)
    v, hh = projection_sweep(binary, Ms, threads=threads)
    return rule_a(hh, v, N, step)


def get_angle_with_projections(bgr, max_angle, step, resize_scale, threads=1):
    """projection::get_angle_with_projections (projection.rs:17-194), single-thread
    semantics (the threads>1 branch of the reference has a unit bug)."""
    binary = preprocess_projection(bgr, resize_scale)
    h, w = binary.shape
    N, angles, Ms = sweep_matrices(w, h, max_angle, step, 1.0)
    v, hh = projection_sweep(binary, Ms, threads=threads)
    return rule_b(v, hh, N, step)


# --------------------------------------------------------------------------- edges / lines
def canny(img, lo, hi):
    """cv::Canny(lo, hi, 3, false) (hough.rs:27, omr.rs:239, fft.rs:164-171)."""
    img, w, h, ch = _img(img)
    out = np.empty((h, w), np.uint8)
    lib().orc_canny(_p(img), w, h, w * ch, ch, C.c_double(lo), C.c_double(hi), _p(out), w)
    return out


def hough_dims(w, h, rho=1.0, theta=math.pi / 180.0):
    na, nr = C.c_int(), C.c_int()
    lib().orc_hough_dims(w, h, C.c_double(rho), C.c_double(theta), C.byref(na), C.byref(nr))
    return na.value, nr.value


def hough_accumulate(edges, rho=1.0, theta=math.pi / 180.0):
    """Standard accumulator with the HoughLinesP vote formula."""
    edges, w, h, ch = _img(edges)
    na, nr = hough_dims(w, h, rho, theta)
    acc = np.zeros((na, nr), np.int32)
    lib().orc_hough_accumulate(_p(edges), w, h, w, C.c_double(rho), C.c_double(theta), _p(acc, C.c_int32))
    return acc


def hough_peaks(accum, threshold):
    """cv::HoughLines' local-maximum rule + ordering over accum [numangle, numrho]
    -> int32 [n, 3] {theta index, rho index, count}."""
    accum = np.ascontiguousarray(accum, np.int32)
    na, nr = accum.shape
    cap = 1 << 12
    while True:
        peaks = np.zeros((cap, 3), np.int32)
        n = lib().orc_hough_peaks(_p(accum, C.c_int32), na, nr, int(threshold), _p(peaks, C.c_int32), cap)
        if n <= cap:
            return peaks[:n].copy()
        cap = n


def scale_self(img, scale):
    """TransformableMatrix::scale_self (transfer.rs:66-91), shrinking side only."""
    if scale == 1.0:
        return img
    assert scale < 1.0, "INTER_LINEAR enlargement is outside the restated path"
    h, w = img.shape[:2]
    return resize_area(img, int(float(w) * scale), int(float(h) * scale))


def shrink_to(img, max_w, max_h):
    """TransformableMatrix::shrink_to (transfer.rs:93-126) -> (image, scale used)."""
    h, w = img.shape[:2]
    s = projection_scale(w, h, max_w, max_h)
    if s >= 1.0:
        return img, 1.0
    return scale_self(img, s), s


def hough_lines_p(edges, threshold, min_len, max_gap, rho=1.0, theta=math.pi / 180.0):
    """cv::HoughLinesP (hough.rs:31-42, omr.rs:245-253, fft.rs:183-191) -> int32 [n,4]."""
    edges, w, h, ch = _img(edges)
    cap = 1 << 16
    while True:
        lines = np.zeros((cap, 4), np.int32)
        n = lib().orc_hough_lines_p(_p(edges), w, h, w, C.c_double(rho), C.c_double(theta), int(threshold),
                                    C.c_double(min_len), C.c_double(max_gap), _p(lines, C.c_int32), cap)
        if n <= cap:
            return lines[:n].copy()
        cap = n


def lines_angle_omr(lines):
    """omr.rs:257-301 -> OmrResult (raises on no lines, like the reference's panic)."""
    lines = np.ascontiguousarray(lines, np.int32).reshape(-1, 4)
    n = len(lines)
    if n == 0:
        raise IndexError("no line segments (reference panics at omr.rs:272)")
    cand = np.zeros(n, np.float64)
    angle, status, nc = C.c_double(), C.c_int(), C.c_int()
    lib().orc_lines_angle_omr(_p(lines, C.c_int32), n, C.byref(angle), C.byref(status),
                              _p(cand, C.c_double), n, C.byref(nc))
    return OmrResult(angle.value, status.value, list(cand[:nc.value]))


def lines_angle_hough(lines):
    """hough.rs:48-89."""
    lines = np.ascontiguousarray(lines, np.int32).reshape(-1, 4)
    if len(lines) == 0:
        raise IndexError("no line segments (reference panics at hough.rs:74)")
    angle = C.c_double()
    lib().orc_lines_angle_hough(_p(lines, C.c_int32), len(lines), C.byref(angle))
    return angle.value


def lines_angle_fft(lines):
    """fft.rs:195-246 (with the index bug at fft.rs:231)."""
    lines = np.ascontiguousarray(lines, np.int32).reshape(-1, 4)
    return lib().orc_lines_angle_fft(_p(lines, C.c_int32), len(lines))


def get_result_from_edges_detection(img, min_len, max_gap):
    """omr::get_result_from_edges_detection (omr.rs:231-301)."""
    edges = canny(img, 50.0, 150.0)
    return lines_angle_omr(hough_lines_p(edges, 0, min_len, max_gap))


def get_angle_with_hough(gray_img, min_len, max_gap):
    """hough::get_angle_with_hough (hough.rs:17-100), minus the debug image."""
    edges = canny(gray_img, 50.0, 150.0)
    return lines_angle_hough(hough_lines_p(edges, 0, min_len, max_gap))


# --------------------------------------------------------------------------- FFT image
def fft_image(gray_img):
    """fft::get_fft_image (fft.rs:124-141) -> (mag8, log8).

    float64 numpy FFT with every convertTo restated as f32(f64(v)*alpha+beta).
    cv::dft itself is float32 and not bit-stable between its own IPP/non-IPP
    paths: parity on log8 is +-1 LSB on <= 1e-4 of the pixels.
    """
    g = np.ascontiguousarray(gray_img, np.uint8)
    h, w = g.shape
    f = (g.astype(np.float64) * (1.0 / 255.0)).astype(np.float32)                  # fft.rs:127-130
    F = np.fft.fft2(f.astype(np.float64)) / float(w * h)                            # fft.rs:50-56 (DFT_SCALE)
    re = F.real.astype(np.float32)
    im = F.imag.astype(np.float32)

    def shift(p):                                                                   # fft.rs:67-86
        q = p.copy()
        cx, cy = w // 2, h // 2
        q[0:cy, 0:cx] = p[cy:2 * cy, cx:2 * cx]
        q[cy:2 * cy, cx:2 * cx] = p[0:cy, 0:cx]
        q[0:cy, cx:2 * cx] = p[cy:2 * cy, 0:cx]
        q[cy:2 * cy, 0:cx] = p[0:cy, cx:2 * cx]
        return q

    re, im = shift(re), shift(im)
    mag = np.sqrt(re.astype(np.float32) ** 2 + im.astype(np.float32) ** 2).astype(np.float32)  # fft.rs:110

    def correction(a):                                                              # fft.rs:88-106
        mn, mx = float(a.min()), float(a.max())
        b = (a.astype(np.float64) * 1.0 + (-mn)).astype(np.float32)
        return (b.astype(np.float64) * (1.0 / (mx - mn)) + 0.0).astype(np.float32)

    def to_u8(a):                                                                   # convertTo(CV_8UC1, 255)
        return np.clip(np.rint(a.astype(np.float64) * 255.0), 0, 255).astype(np.uint8)

    m = (correction(mag).astype(np.float64) * 255.0).astype(np.float32)             # fft.rs:112-113
    mag8 = to_u8(m)                                                                 # fft.rs:134-135
    lg = (m.astype(np.float64) * 1.0 + (1.0 / 255.0)).astype(np.float32)            # fft.rs:118
    lg = np.log(lg.astype(np.float32)).astype(np.float32)                           # fft.rs:119
    log8 = to_u8(correction(lg))                                                    # fft.rs:120,137-138
    return mag8, log8


def get_angle_with_fft(gray_img, canny1, canny2, min_len, max_gap):
    """fft::get_angle_with_fft (fft.rs:145-256), minus the debug image."""
    _, log8 = fft_image(gray_img)
    edges = canny(log8, canny1, canny2)
    return lines_angle_fft(hough_lines_p(edges, 100, min_len, max_gap))


def get_result_from_fourier_transform(bgr, canny1, canny2, min_len, max_gap):
    """omr::get_result_from_fourier_transform (omr.rs:304-337): Canny twice, thr 0."""
    _, log8 = fft_image(gray(bgr))
    edges = canny(log8, canny1, canny2)
    return get_result_from_edges_detection(edges, min_len, max_gap)


# --------------------------------------------------------------------------- one-call pipeline
def reconcile(p: OmrResult, e_angle):
    """omr.rs:352-403."""
    cand = np.ascontiguousarray(p.candidates, np.float64)
    angle, need = C.c_double(), C.c_int()
    lib().orc_reconcile(p.status, C.c_double(p.angle), _p(cand, C.c_double) if len(cand) else None,
                        len(cand), C.c_double(e_angle), C.byref(angle), C.byref(need))
    return angle.value, bool(need.value)


def correct_default(bgr, max_angle, step, max_w, max_h, hough_min_len, hough_max_gap, threads=1):
    """omr::correct_default (omr.rs:339-452) on memory buffers (imread/imwrite stay
    with the caller) -> (angle, need_check, deskewed BGR)."""
    p = get_result_from_projection(bgr, max_angle, step, max_w, max_h, threads=threads)
    if p.status == BELIEVED:
        angle, need = p.angle, False
    else:
        e = get_result_from_edges_detection(bgr, hough_min_len, hough_max_gap)
        angle, need = reconcile(p, e.angle)
    out = rotate_mat(bgr, angle, 1.0, INTER_NEAREST, WHITE, CLIP_CONTAIN)            # omr.rs:408-445
    return angle, need, out


# --------------------------------------------------------------------------- N2: imwrite(JPEG, q)
def jpeg_encode(bgr, quality=100) -> bytes:
    """TransformableMatrix::im_write(path, JPEG, 100) (transfer.rs:162-179, omr.rs:447-448) ->
    the bytes cv::imwrite / cv::imencode produce for a BGR image."""
    bgr, w, h, ch = _img(bgr)
    assert ch == 3
    lib().orc_jpeg_encode.restype = C.c_long
    cap = 4096 + 4 * w * h
    out = np.empty(cap, np.uint8)
    n = lib().orc_jpeg_encode(_p(bgr), w, h, w * 3, int(quality), _p(out), C.c_long(cap))
    assert n <= cap
    return out[:n].tobytes()


# --------------------------------------------------------------------------- N2: imread(JPEG)
def jpeg_decode(data) -> np.ndarray:
    """imgcodecs::imread(path, IMREAD_COLOR) (omr.rs:349, transfer.rs:43) of a baseline JPEG file's
    bytes -> BGR image.  Raises ValueError for corrupt (-1) or unsupported (-2) files."""
    buf = np.frombuffer(bytes(data), np.uint8)
    w, h, nc = C.c_int(0), C.c_int(0), C.c_int(0)
    rc = lib().orc_jpeg_info(_p(buf), C.c_long(buf.size), C.byref(w), C.byref(h), C.byref(nc))
    if rc:
        raise ValueError(f"jpeg: header rejected ({rc})")
    out = np.empty((h.value, w.value, 3), np.uint8)
    rc = lib().orc_jpeg_decode(_p(buf), C.c_long(buf.size), _p(out), C.c_long(w.value * 3))
    if rc:
        raise ValueError(f"jpeg: decode failed ({rc})")
    return out
