"""Golden vectors minted from the real OpenCV (tests/golden/make_golden.py, cv2 4.13 driven
with the reference's call sequence on cards of the reference's own dataset).

The same checks run twice: against the CPU oracle (no GPU needed -- this is what pins the
oracle where cv2 is absent) and against the CUDA path through the C ABI (`-m gpu`)."""
import glob
import hashlib
import os

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
FILES = sorted(glob.glob(os.path.join(HERE, "golden", "*.npz")))
WHITE = (255, 255, 255, 0)


def _sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def _unpack(bits, shape):
    n = int(shape[0]) * int(shape[1])
    return np.unpackbits(bits)[:n].reshape(int(shape[0]), int(shape[1])).astype(bool)


class OracleImpl:
    """Adapter: the oracle behind the names the checks use."""

    def __init__(self, orc):
        self.o = orc

    def preprocess_omr(self, bgr, mw, mh):
        return self.o.preprocess_omr(bgr, mw, mh)

    def preprocess_projection(self, bgr, s):
        return self.o.preprocess_projection(bgr, s)

    def sweep(self, b, Ms):
        v, h, rows, cols = self.o.projection_sweep(b, Ms, want_counts=True, threads=4)
        return v, h, rows, cols

    def rule_a(self, h, v, N, step):
        r = self.o.rule_a(h, v, N, step)
        return r.angle, r.status, r.candidates

    def rule_b(self, v, h, N, step):
        return self.o.rule_b(v, h, N, step)

    def canny(self, img, lo, hi):
        return self.o.canny(img, lo, hi)

    def hough_lines_p(self, e, thr, ml, mg):
        return self.o.hough_lines_p(e, thr, ml, mg)

    def lines_omr(self, L):
        r = self.o.lines_angle_omr(L)
        return r.angle, r.status, r.candidates

    def lines_hough(self, L):
        return self.o.lines_angle_hough(L)

    def lines_fft(self, L):
        return self.o.lines_angle_fft(L)

    def fft_log8(self, g):
        return self.o.fft_image(g)[1]

    def rotate_mat(self, img, ang, flags, clip):
        return self.o.rotate_mat(img, ang, 1.0, flags, WHITE, clip)


class CudaImpl:
    def __init__(self, ctx, api):
        self.c, self.api = ctx, api

    def preprocess_omr(self, bgr, mw, mh):
        return self.c.preprocess(bgr, self.api.PRE_OMR, mw, mh)

    def preprocess_projection(self, bgr, s):
        return self.c.preprocess(bgr, self.api.PRE_PROJECTION, resize_scale=s)[0]

    def sweep(self, b, Ms):
        v, h = self.c.projection_sweep(b, Ms)
        rows, cols = zip(*[self.c.projection_counts(b, Ms[i]) for i in range(0, len(Ms), 7)])
        full_r = np.zeros((len(Ms), b.shape[0]), np.int32)
        full_c = np.zeros((len(Ms), b.shape[1]), np.int32)
        full_r[::7], full_c[::7] = np.stack(rows), np.stack(cols)
        return v, h, full_r, full_c

    def rule_a(self, h, v, N, step):
        r = self.api.rule_a(h, v, N, step)
        return r.angle, r.status, r.candidates

    def rule_b(self, v, h, N, step):
        return self.api.rule_b(v, h, N, step)

    def canny(self, img, lo, hi):
        return self.c.canny(img, lo, hi)

    def hough_lines_p(self, e, thr, ml, mg):
        return self.c.hough_lines_p(e, thr, ml, mg)

    def lines_omr(self, L):
        r = self.api.lines_angle(L, self.api.LINES_OMR)
        return r.angle, r.status, r.candidates

    def lines_hough(self, L):
        return self.api.lines_angle(L, self.api.LINES_HOUGH).angle

    def lines_fft(self, L):
        return self.api.lines_angle(L, self.api.LINES_FFT).angle

    def fft_log8(self, g):
        return self.c.fft_image(g)[1]

    def rotate_mat(self, img, ang, flags, clip):
        return self.c.rotate_mat(img, ang, 1.0, flags, WHITE, clip)


def _sweep_mats(orc_or_api, w, h, max_angle, step, scale):
    return orc_or_api.sweep_matrices(w, h, int(max_angle), float(step), float(scale))


def _check(impl, geo, path):
    G = np.load(path, allow_pickle=False)
    gray_in = G["gray_in"]
    bgr = np.ascontiguousarray(np.repeat(gray_in[:, :, None], 3, axis=2))
    # ---- P0 + sweep + rule A
    for tag in ("def", "full"):
        max_angle, step, mw, mh = G[f"{tag}_params"]
        b, s = impl.preprocess_omr(bgr, int(mw), int(mh))
        want = _unpack(G[f"{tag}_bin"], G[f"{tag}_bin_shape"])
        assert s == float(G[f"{tag}_scale"]) and b.shape == want.shape and np.array_equal(b == 0, want), tag
        N, angles, Ms = _sweep_mats(geo, b.shape[1], b.shape[0], max_angle, step, s)
        v, h, rows, cols = impl.sweep(b, Ms)
        assert np.array_equal(rows[::7], G[f"{tag}_rows"]) and np.array_equal(cols[::7], G[f"{tag}_cols"]), tag
        # north_star: std-devs within 1e-4 relative; the restated evaluation order gives equality
        assert np.allclose(v, G[f"{tag}_v_sd"], rtol=1e-12, atol=0) and np.allclose(h, G[f"{tag}_h_sd"], rtol=1e-12, atol=0)
        assert np.array_equal(v, G[f"{tag}_v_sd"]) and np.array_equal(h, G[f"{tag}_h_sd"]), tag
        a, st, cand = impl.rule_a(h, v, N, float(step))
        assert (a, st) == (float(G[f"{tag}_rule_a"][0]), int(G[f"{tag}_rule_a"][1])) and cand == list(G[f"{tag}_rule_a_cand"])
    # ---- P0' + rule B
    b = impl.preprocess_projection(bgr, 0.4)
    assert np.array_equal(b == 0, _unpack(G["proj_bin"], G["proj_bin_shape"]))
    N, angles, Ms = _sweep_mats(geo, b.shape[1], b.shape[0], 45, 0.5, 1.0)
    v, h, _, _ = impl.sweep(b, Ms)
    assert np.array_equal(v, G["proj_v_sd"]) and np.array_equal(h, G["proj_h_sd"])
    assert impl.rule_b(v, h, N, 0.5) == float(G["proj_rule_b"])
    # ---- H1 / H2 / H3
    e_gray = impl.canny(gray_in, 50.0, 150.0)
    assert np.array_equal(e_gray != 0, _unpack(G["canny_gray"], gray_in.shape))
    assert np.array_equal(impl.canny(bgr, 50.0, 150.0) != 0, _unpack(G["canny_bgr"], gray_in.shape))
    for key in [k for k in G.files if k.startswith("lines_") and k.count("_") == 3]:
        _, thr, ml, mg = key.split("_")
        L = impl.hough_lines_p(e_gray, int(thr), float(ml), float(mg))
        assert L.shape == G[key].shape and np.array_equal(L, G[key]), key
        if len(L):
            a, st, cand = impl.lines_omr(L)
            assert (a, st) == (float(G[key + "_omr"][0]), int(G[key + "_omr"][1])) and cand == list(G[key + "_omr_cand"])
            assert impl.lines_hough(L) == float(G[key + "_hough"])
            assert impl.lines_fft(L) == float(G[key + "_fft"])
    # ---- F1: cv::dft is float32 (and not bit-stable with itself): +-1 LSB on <= 1e-4 of the pixels
    log8 = impl.fft_log8(gray_in)
    d = np.abs(log8.astype(np.int16) - G["fft_log8"].astype(np.int16))
    assert d.max() <= 1 and np.count_nonzero(d) <= 1e-4 * d.size, (int(d.max()), int(np.count_nonzero(d)))
    # Canny + HoughLinesP(thr 100) on the GOLDEN log image (isolates H1/H2 from the +-1 LSB)
    e_fft = impl.canny(G["fft_log8"], 125.0, 150.0)
    assert np.array_equal(e_fft != 0, _unpack(G["fft_canny"], gray_in.shape))
    assert np.array_equal(impl.hough_lines_p(e_fft, 100, 60.0, 8.0), G["fft_lines"])
    # ---- W1
    ang = float(G["warp_angle"])
    for flags in (0, 1):
        for clip in (0, 1):
            r = impl.rotate_mat(bgr, ang, flags, clip)
            assert list(r.shape) == list(G[f"warp_{flags}_{clip}_shape"])
            assert _sha(r) == str(G[f"warp_{flags}_{clip}_sha"]), (flags, clip)


def test_golden_files_exist():
    assert len(FILES) >= 2


@pytest.mark.parametrize("path", FILES, ids=[os.path.basename(f) for f in FILES])
def test_oracle_matches_golden(orc, path):
    _check(OracleImpl(orc), orc, path)


@pytest.mark.gpu
@pytest.mark.parametrize("path", FILES, ids=[os.path.basename(f) for f in FILES])
def test_cuda_matches_golden(path):
    from omr_img_corrector_b200 import api
    with api.Context(0) as ctx:
        _check(CudaImpl(ctx, api), api, path)


# ----------------------------------------------------------------------------- N2: JPEG files
JPEGS = sorted(glob.glob(os.path.join(HERE, "golden", "jpeg", "*.jpg")))


def _jpeg_expected():
    return np.load(os.path.join(HERE, "golden", "jpeg", "expected_pixels.npz"))


def test_jpeg_goldens_exist():
    assert len(JPEGS) >= 12


@pytest.mark.parametrize("path", JPEGS, ids=[os.path.basename(f) for f in JPEGS])
def test_oracle_decodes_golden_jpeg(orc, path):
    """cv2.imdecode's pixels of cv2-written files (tools/make_jpeg_goldens.py), bit for bit."""
    want = _jpeg_expected()[os.path.basename(path)[:-4]]
    assert np.array_equal(orc.jpeg_decode(open(path, "rb").read()), want)


@pytest.mark.parametrize("path", [p for p in JPEGS if "420_q100" in p or "420_q75" in p],
                         ids=lambda p: os.path.basename(p))
def test_oracle_reencodes_golden_jpeg(orc, path):
    """imwrite side: encoding the decoded pixels at the file's quality reproduces cv2's own encoding of them."""
    import cv2
    px = _jpeg_expected()[os.path.basename(path)[:-4]]
    q = 100 if "q100" in path else 75
    assert orc.jpeg_encode(px, q) == bytes(cv2.imencode(".jpg", px, [cv2.IMWRITE_JPEG_QUALITY, q])[1])


@pytest.mark.gpu
@pytest.mark.parametrize("path", JPEGS, ids=[os.path.basename(f) for f in JPEGS])
def test_cuda_decodes_golden_jpeg(path):
    from omr_img_corrector_b200 import api
    want = _jpeg_expected()[os.path.basename(path)[:-4]]
    with api.Context(0) as ctx:
        assert np.array_equal(ctx.jpeg_decode(open(path, "rb").read()), want)
