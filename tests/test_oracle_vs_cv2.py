"""Pins the CPU oracle against the real OpenCV build importable in this image.

The reference delegates all pixel arithmetic to OpenCV (third-party, not in
the reference tree) and holds no golden vectors of its own (SURVEY.md section 4,
8(c)); these tests are what makes the oracle trustworthy.  They skip when cv2
is not importable; the committed fixtures in tests/golden/ (minted from cv2)
cover that case.
"""
import math

import numpy as np
import pytest

cv2 = pytest.importorskip("cv2")

from omr_img_corrector_b200 import synth  # noqa: E402

WHITE = (255, 255, 255, 0)


def _cv_rot(w, h, angle, scale):
    return cv2.getRotationMatrix2D((float(np.float32(w) / np.float32(2)), float(np.float32(h) / np.float32(2))),
                                   angle, scale)


@pytest.fixture(scope="module")
def small_card(orc):
    g, a = synth.base_card(3)
    g = np.ascontiguousarray(g[200:1350, 100:1340])            # 1150 x 1240 crop, dataset-like
    bgr = synth.to_bgr(g)
    return orc.rotate_mat(bgr, -a, 1.0, 1, WHITE, 0), a


def test_rotation_matrix_matches_cv2(orc):
    for (w, h) in [(248, 230), (2480, 3508), (1237, 1300)]:
        for ang in [-15.0, -3.7, 0.1, 3 * 0.1, 44.8, -45.0]:
            for s in [1.0, 0.2, 0.1769]:
                M = orc.rotation_matrix_2d(np.float32(w) / np.float32(2), np.float32(h) / np.float32(2), ang, s)
                assert np.array_equal(M.reshape(2, 3), _cv_rot(w, h, ang, s))


@pytest.mark.parametrize("flags", [0, 1])
@pytest.mark.parametrize("ch", [1, 3])
def test_warp_affine_bit_exact(orc, flags, ch):
    rng = np.random.default_rng(7 + flags + ch)
    shape = (301, 417) if ch == 1 else (301, 417, 3)
    img = rng.integers(0, 256, shape, dtype=np.uint8)
    h, w = shape[:2]
    for ang in [-15.0, -3.7, 0.1, 7.3, 44.8, -90.0, 180.0]:
        for s in [1.0, 0.2]:
            M = _cv_rot(w, h, ang, s)
            for ds in [(w, h), (w + 37, h + 11)]:
                ref = cv2.warpAffine(img, M, ds, flags=flags, borderMode=cv2.BORDER_CONSTANT, borderValue=WHITE)
                got = orc.warp_affine(img, M, ds, flags, WHITE)
                assert np.array_equal(ref, got), (ang, s, ds)


def test_rotate_mat_contain_geometry(orc):
    rng = np.random.default_rng(11)
    img = rng.integers(0, 256, (230, 248, 3), dtype=np.uint8)
    for ang in [-14.9, -2.8, 0.0, 3.7, 15.0]:
        a = ang * math.pi / 180.0
        rw = math.ceil(230 * abs(math.sin(a)) + 248 * abs(math.cos(a)))
        rh = math.ceil(248 * abs(math.sin(a)) + 230 * abs(math.cos(a)))
        M = cv2.getRotationMatrix2D((float(np.float32(math.ceil(rw / 2))), float(np.float32(math.ceil(rh / 2)))), ang, 1.0)
        M[0, 2] += math.ceil((rw - 248) / 2)
        M[1, 2] += math.ceil((rh - 230) / 2)
        for flags in (0, 1):
            ref = cv2.warpAffine(img, M, (int(rw), int(rh)), flags=flags, borderMode=cv2.BORDER_CONSTANT,
                                 borderValue=WHITE)
            got = orc.rotate_mat(img, ang, 1.0, flags, WHITE, orc.CLIP_CONTAIN)
            assert got.shape == ref.shape and np.array_equal(ref, got)


def test_preprocess_pieces(orc):
    rng = np.random.default_rng(5)
    bgr = rng.integers(0, 256, (233, 341, 3), dtype=np.uint8)
    g = cv2.cvtColor(bgr, cv2.COLOR_RGB2GRAY)
    assert np.array_equal(g, orc.gray(bgr))
    k = cv2.getStructuringElement(cv2.MORPH_ELLIPSE, (3, 3), (-1, -1))
    e = cv2.erode(g, k, anchor=(-1, -1), iterations=3, borderType=cv2.BORDER_CONSTANT)
    assert np.array_equal(e, orc.erode3(g, 3))
    _, t = cv2.threshold(g, 127, 255, cv2.THRESH_BINARY)
    assert np.array_equal(t, orc.threshold(g))


@pytest.mark.parametrize("shape,dsize", [
    ((1150, 1240), (248, 230)),      # dataset default: exact 5x
    ((1300, 1237), (218, 230)),      # SCN cards: fractional 0.1769
    ((1300, 1236), (218, 230)),
    ((600, 400), (200, 300)),        # 2x2 special path
    ((601, 403), (133, 77)),
    ((301, 417), (417, 301)),        # identity
    ((3508 // 4, 2480 // 4), (248, 350)),
])
def test_resize_area(orc, shape, dsize):
    rng = np.random.default_rng(sum(shape))
    for ch in (1, 3):
        img = rng.integers(0, 256, shape if ch == 1 else shape + (3,), dtype=np.uint8)
        ref = cv2.resize(img, dsize, interpolation=cv2.INTER_AREA)
        got = orc.resize_area(img, dsize[0], dsize[1])
        assert np.array_equal(ref, got)


def test_preprocess_omr_default_operating_point(orc, small_card):
    bgr, _ = small_card
    b, s = orc.preprocess_omr(bgr, 248, 230)
    h, w = bgr.shape[:2]
    g = cv2.cvtColor(bgr, cv2.COLOR_RGB2GRAY)
    k = cv2.getStructuringElement(cv2.MORPH_ELLIPSE, (3, 3), (-1, -1))
    e = cv2.erode(g, k, anchor=(-1, -1), iterations=3, borderType=cv2.BORDER_CONSTANT)
    r = cv2.resize(e, (int(w * s), int(h * s)), interpolation=cv2.INTER_AREA)
    _, t = cv2.threshold(r, 127, 255, cv2.THRESH_BINARY)
    assert b.shape == (230, 248) and np.array_equal(b, t)


def test_projection_sweep_matches_cv2_sequence(orc, small_card):
    """The hot loop restated with cv2 calls (omr.rs:153-208) vs the C oracle."""
    bgr, a = small_card
    b, s = orc.preprocess_omr(bgr, 248, 230)
    h, w = b.shape
    N, angles, Ms = orc.sweep_matrices(w, h, 45, 0.2, s)
    assert N == 225 and len(angles) == 450
    v, hh, rows, cols = orc.projection_sweep(b, Ms, want_counts=True)
    for i in range(0, 450, 7):
        M = _cv_rot(w, h, angles[i], s)
        assert np.array_equal(M.reshape(6), Ms[i])
        rot = cv2.warpAffine(b, M, (w, h), flags=cv2.WARP_POLAR_LINEAR, borderMode=cv2.BORDER_CONSTANT,
                             borderValue=WHITE)
        r = (rot == 0).sum(axis=1).astype(np.float64)
        c = (rot == 0).sum(axis=0).astype(np.float64)
        assert np.array_equal(r, rows[i]) and np.array_equal(c, cols[i])
        # calculate.rs in Python floats, with ** as Rust's powf
        def sd(vec):
            m = vec[0]
            for x in vec[1:]:
                m = m + x
            m /= len(vec)
            acc = (vec[0] - m) ** 2.0
            for x in vec[1:]:
                acc = acc + (x - m) ** 2.0
            return (acc / len(vec)) ** 0.5
        assert abs(sd(list(r)) - hh[i]) <= 1e-12 * max(1.0, hh[i])
        assert abs(sd(list(c)) - v[i]) <= 1e-12 * max(1.0, v[i])


@pytest.mark.parametrize("ch", [1, 3])
def test_canny_bit_exact(orc, small_card, ch):
    bgr, _ = small_card
    rng = np.random.default_rng(2)
    noise = rng.integers(0, 256, (211, 173, 3), dtype=np.uint8)
    cases = [bgr, noise, cv2.GaussianBlur(noise, (5, 5), 0)]
    for img in cases:
        src = img if ch == 3 else np.ascontiguousarray(img[:, :, 0])
        for lo, hi in [(50.0, 150.0), (125.0, 150.0), (150.0, 75.0)]:
            ref = cv2.Canny(src, lo, hi, apertureSize=3, L2gradient=False)
            got = orc.canny(src, lo, hi)
            assert np.array_equal(ref, got), (ch, lo, hi)


def _cv_lines(edges, thr, ml, mg):
    l = cv2.HoughLinesP(edges, 1.0, math.pi / 180.0, thr, minLineLength=ml, maxLineGap=mg)
    return np.zeros((0, 4), np.int32) if l is None else l.reshape(-1, 4).astype(np.int32)


def test_hough_lines_p_line_for_line(orc, small_card):
    bgr, _ = small_card
    edges_bgr = cv2.Canny(bgr, 50.0, 150.0)
    edges_small = cv2.Canny(cv2.resize(bgr, (310, 288), interpolation=cv2.INTER_AREA), 50.0, 150.0)
    for edges in (edges_bgr, edges_small):
        for thr, ml, mg in [(0, 125.0, 15.0), (0, 150.0, 50.0), (100, 125.0, 15.0), (30, 40.0, 5.0), (0, 20.5, 3.5)]:
            ref = _cv_lines(edges, thr, ml, mg)
            got = orc.hough_lines_p(edges, thr, ml, mg)
            assert ref.shape == got.shape and np.array_equal(ref, got), (thr, ml, mg, len(ref), len(got))


def test_fft_image_within_one_lsb(orc, small_card):
    bgr, _ = small_card
    g = cv2.cvtColor(bgr, cv2.COLOR_RGB2GRAY)
    # the reference's call sequence (fft.rs:42-141) through cv2
    src = (g.astype(np.float64) * (1.0 / 255.0)).astype(np.float32)
    cplx = cv2.merge([src, np.zeros_like(src)])
    d = cv2.dft(cplx, flags=cv2.DFT_COMPLEX_OUTPUT | cv2.DFT_COMPLEX_INPUT | cv2.DFT_SCALE)
    re, im = cv2.split(d)

    def shift(p):
        q = p.copy()
        h, w = p.shape
        cx, cy = w // 2, h // 2
        q[0:cy, 0:cx] = p[cy:2 * cy, cx:2 * cx]
        q[cy:2 * cy, cx:2 * cx] = p[0:cy, 0:cx]
        q[0:cy, cx:2 * cx] = p[cy:2 * cy, 0:cx]
        q[cy:2 * cy, 0:cx] = p[0:cy, cx:2 * cx]
        return q

    mag = cv2.magnitude(shift(re), shift(im))

    def conv(a, alpha, beta):
        return (a.astype(np.float64) * alpha + beta).astype(np.float32)

    def correction(a):
        mn, mx, _, _ = cv2.minMaxLoc(a)
        return conv(conv(a, 1.0, -mn), 1.0 / (mx - mn), 0.0)

    m = conv(correction(mag), 255.0, 0.0)
    lg = cv2.log(conv(m, 1.0, 1.0 / 255.0))
    log8 = np.clip(np.rint(correction(lg).astype(np.float64) * 255.0), 0, 255).astype(np.uint8)
    _, got = orc.fft_image(g)
    diff = np.abs(log8.astype(np.int32) - got.astype(np.int32))
    assert diff.max() <= 1
    assert (diff != 0).mean() <= 1e-3


# ----------------------------------------------------------------------------- N2: imwrite(JPEG, q)
@pytest.mark.parametrize("shape", [(37, 53), (64, 48), (16, 16), (95, 130), (1, 1), (9, 7), (24, 40), (33, 17), (230, 248)])
def test_jpeg_encode_bytes_equal_cv2(orc, shape):
    """omr.rs:447-448: the deskewed card is written with imwrite(JPEG, quality 100).  The restatement
    must produce cv2's file byte for byte (also at a few other qualities)."""
    rng = np.random.default_rng(shape[0] * 1000 + shape[1])
    img = (rng.random(shape + (3,)) * 255).astype(np.uint8)
    img[shape[0] // 4:shape[0] // 2, shape[1] // 5:shape[1] // 2] = (10, 200, 90)
    smooth = cv2.GaussianBlur(img, (7, 7), 2.0)
    for im in (img, smooth, np.full(shape + (3,), 255, np.uint8)):
        for q in (100, 95, 75, 30):
            ref = bytes(cv2.imencode(".jpg", im, [cv2.IMWRITE_JPEG_QUALITY, q])[1])
            assert orc.jpeg_encode(im, q) == ref, (shape, q)


# ----------------------------------------------------------------------------- N2: imread(JPEG)
_JPEG_VARIANTS = {
    "420": [],
    "444": [cv2.IMWRITE_JPEG_SAMPLING_FACTOR, cv2.IMWRITE_JPEG_SAMPLING_FACTOR_444],
    "422": [cv2.IMWRITE_JPEG_SAMPLING_FACTOR, cv2.IMWRITE_JPEG_SAMPLING_FACTOR_422],
    "optimised-tables": [cv2.IMWRITE_JPEG_OPTIMIZE, 1],
    "restart-3": [cv2.IMWRITE_JPEG_RST_INTERVAL, 3],
}


@pytest.mark.parametrize("shape", [(37, 53), (64, 48), (16, 16), (1, 1), (5, 4), (9, 2), (33, 17), (230, 248)])
def test_jpeg_decode_equals_cv2(orc, shape):
    """omr.rs:349: the card is read with imread(path, IMREAD_COLOR).  The restatement of libjpeg's
    baseline decompressor must give cv2's pixels bit for bit: 4:2:0 / 4:2:2 / 4:4:4 / grey files,
    optimised Huffman tables, restart intervals, several qualities."""
    rng = np.random.default_rng(shape[0] * 1000 + shape[1])
    img = (rng.random(shape + (3,)) * 255).astype(np.uint8)
    img[shape[0] // 4:shape[0] // 2, shape[1] // 5:shape[1] // 2] = (10, 200, 90)
    smooth = cv2.GaussianBlur(img, (7, 7), 2.0)
    for im in (img, smooth, np.full(shape + (3,), 255, np.uint8)):
        for q in (100, 90, 30):
            for name, extra in _JPEG_VARIANTS.items():
                enc = cv2.imencode(".jpg", im, [cv2.IMWRITE_JPEG_QUALITY, q] + extra)[1]
                assert np.array_equal(orc.jpeg_decode(enc.tobytes()), cv2.imdecode(enc, cv2.IMREAD_COLOR)), (shape, q, name)
        enc = cv2.imencode(".jpg", cv2.cvtColor(im, cv2.COLOR_BGR2GRAY), [cv2.IMWRITE_JPEG_QUALITY, 90])[1]
        assert np.array_equal(orc.jpeg_decode(enc.tobytes()), cv2.imdecode(enc, cv2.IMREAD_COLOR)), (shape, "grey")


def test_jpeg_decode_rejects_what_it_does_not_cover(orc):
    img = np.zeros((16, 16, 3), np.uint8)
    prog = cv2.imencode(".jpg", img, [cv2.IMWRITE_JPEG_PROGRESSIVE, 1])[1].tobytes()
    with pytest.raises(ValueError):
        orc.jpeg_decode(prog)
    with pytest.raises(ValueError):
        orc.jpeg_decode(b"not a jpeg at all")


def _with_exif_orientation(jpeg: bytes, orientation: int) -> bytes:
    """Insert an APP1 Exif segment holding only the Orientation tag behind SOI."""
    tiff = b"II*\x00\x08\x00\x00\x00" + b"\x01\x00" + b"\x12\x01\x03\x00\x01\x00\x00\x00" + bytes([orientation, 0, 0, 0]) + b"\x00\x00\x00\x00"
    body = b"Exif\x00\x00" + tiff
    return jpeg[:2] + b"\xff\xe1" + (len(body) + 2).to_bytes(2, "big") + body + jpeg[2:]


def test_jpeg_decode_exif_orientation(orc):
    """cv::imread turns the image by the Exif Orientation tag; the restatement (and the CUDA decoder,
    which shares the rule) accepts upright files only and refuses the others instead of returning
    pixels imread would not."""
    img = np.zeros((24, 40, 3), np.uint8)
    img[:8, :10] = (0, 0, 255)
    enc = cv2.imencode(".jpg", img, [cv2.IMWRITE_JPEG_QUALITY, 95])[1].tobytes()
    upright = _with_exif_orientation(enc, 1)
    assert np.array_equal(orc.jpeg_decode(upright), cv2.imdecode(np.frombuffer(upright, np.uint8), cv2.IMREAD_COLOR))
    turned = _with_exif_orientation(enc, 6)
    assert cv2.imdecode(np.frombuffer(turned, np.uint8), cv2.IMREAD_COLOR).shape[:2] == (40, 24)   # cv2 does rotate
    with pytest.raises(ValueError):
        orc.jpeg_decode(turned)


def _standard_accumulator(edges, rho=1.0, theta=math.pi / 180.0):
    """cv::HoughLinesStandard's stage 1 restated in numpy: its trig table accumulates the angle in
    float32 (createTrigTable), unlike HoughLinesP's (float)(n * theta)."""
    h, w = edges.shape
    numangle = int(math.floor(math.pi / np.float32(theta))) + 1
    if numangle > 1 and abs(math.pi - (numangle - 1) * float(np.float32(theta))) < float(np.float32(theta)) / 2:
        numangle -= 1
    numrho = int(round(((w + h) * 2 + 1) / rho))
    irho = np.float32(1.0) / np.float32(rho)
    ang = np.float32(0.0)
    tab_sin, tab_cos = np.zeros(numangle, np.float32), np.zeros(numangle, np.float32)
    for n in range(numangle):
        tab_sin[n] = np.float32(math.sin(float(ang)) * float(irho))
        tab_cos[n] = np.float32(math.cos(float(ang)) * float(irho))
        ang = np.float32(ang + np.float32(theta))
    ys, xs = np.nonzero(edges)
    acc = np.zeros((numangle, numrho), np.int32)
    for n in range(numangle):
        v = xs.astype(np.float32) * tab_cos[n] + ys.astype(np.float32) * tab_sin[n]
        r = np.rint(v).astype(np.int64) + (numrho - 1) // 2
        np.add.at(acc[n], r, 1)
    return acc, numangle, numrho


@pytest.mark.parametrize("threshold", [40, 80, 150])
def test_hough_peaks_match_cv2_hough_lines(orc, threshold):
    """Pins the peak-extraction oracle (local-maximum rule + ordering) to cv2.HoughLines: the same
    accumulator (restated above with HoughLinesStandard's own trig table) must yield cv2's line list,
    in cv2's order."""
    rng = np.random.default_rng(5)
    img = np.zeros((240, 320), np.uint8)
    cv2.line(img, (10, 20), (300, 200), 255, 1)
    cv2.line(img, (20, 220), (310, 30), 255, 1)
    cv2.line(img, (160, 0), (160, 239), 255, 1)
    cv2.line(img, (0, 120), (319, 121), 255, 1)
    img[rng.random(img.shape) < 0.01] = 255
    ref = cv2.HoughLines(img, 1.0, math.pi / 180.0, threshold)
    ref = np.zeros((0, 2), np.float32) if ref is None else ref.reshape(-1, 2)
    acc, numangle, numrho = _standard_accumulator(img)
    peaks = orc.hough_peaks(acc, threshold)
    got = np.stack([(peaks[:, 1].astype(np.float32) - np.float32((numrho - 1) * 0.5)) * np.float32(1.0),
                    (peaks[:, 0].astype(np.float32) * np.float32(math.pi / 180.0))], axis=1).astype(np.float32)
    assert len(got) == len(ref) and len(ref) > 0
    assert np.array_equal(got, ref)


def test_scale_self_and_shrink_to_match_cv2(orc, small_card):
    """T1: TransformableMatrix::scale_self / shrink_to (transfer.rs:66-126) = cv::resize(INTER_AREA)."""
    bgr, _ = small_card
    h, w = bgr.shape[:2]
    for s in [0.2, 0.5, 0.37]:
        ref = cv2.resize(bgr, (int(w * s), int(h * s)), fx=s, fy=s, interpolation=cv2.INTER_AREA)
        assert np.array_equal(orc.scale_self(bgr, s), ref)
    out, s = orc.shrink_to(bgr, 248, 230)
    assert s == min(248 / w, 230 / h)
    assert np.array_equal(out, cv2.resize(bgr, (int(w * s), int(h * s)), fx=s, fy=s, interpolation=cv2.INTER_AREA))
    same, s1 = orc.shrink_to(bgr, 0, 5000)
    assert s1 == 1.0 and same is bgr                       # never upscales
