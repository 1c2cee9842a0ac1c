"""Host-side mirror of the reference `oics` crate's public functions, over the C ABI.

Images are numpy uint8 arrays ([h, w] or [h, w, 3], C-contiguous rows) or any
object exposing a CUDA pointer (`torch.Tensor` on the GPU, or `DeviceImage`).
Every call goes through liboics_cuda.so; nothing here computes pixels.
"""
from __future__ import annotations

import ctypes as C
import math
import threading
from dataclasses import dataclass, field

import numpy as np

from . import _lib
from ._lib import OicsError, OmrResultC, check

BELIEVED, NEED_CHECK, NOT_A_RESULT = 0, 1, 2          # omr::ResultStatus (omr.rs:41-45)
CLIP_DEFAULT, CLIP_CONTAIN = 0, 1                     # types::RotateClipStrategy (types.rs:7-11)
INTER_NEAREST, INTER_LINEAR = 0, 1
WARP_POLAR_LINEAR = 0                                 # the flag value the reference passes (omr.rs:170)
PRE_OMR, PRE_PROJECTION = 0, 1
LINES_OMR, LINES_HOUGH, LINES_FFT = 0, 1, 2
WHITE = (255, 255, 255, 0)


@dataclass
class OmrResult:
    """omr::OmrResult (omr.rs:46-50)."""
    angle: float
    status: int
    candidates: list = field(default_factory=list)


@dataclass
class DeviceImage:
    """An image already resident on the GPU (raw device address)."""
    ptr: int
    w: int
    h: int
    ch: int = 1
    stride: int = 0

    def __post_init__(self):
        if not self.stride:
            self.stride = self.w * self.ch


def _describe(img):
    """-> (pointer, w, h, ch, stride, keepalive)"""
    if hasattr(img, "get_mat") and hasattr(img, "matrix"):       # transfer.TransformableMatrix: the Mat it owns
        img = img.matrix
    if isinstance(img, DeviceImage):
        return img.ptr, img.w, img.h, img.ch, img.stride, img
    if isinstance(img, np.ndarray):
        if img.dtype != np.uint8 or img.ndim not in (2, 3):
            raise TypeError("expected a uint8 [h,w] or [h,w,ch] array")
        if img.strides[-1] != 1 or (img.ndim == 3 and img.strides[1] != img.shape[2]):
            img = np.ascontiguousarray(img)
        h, w = img.shape[:2]
        ch = 1 if img.ndim == 2 else img.shape[2]
        return img.ctypes.data, w, h, ch, img.strides[0], img
    if hasattr(img, "data_ptr") and hasattr(img, "is_cuda"):     # torch tensor
        if not img.is_cuda:
            return _describe(img.numpy())
        if str(img.dtype) != "torch.uint8" or img.dim() not in (2, 3) or not img.is_contiguous():
            raise TypeError("expected a contiguous uint8 CUDA tensor [h,w] or [h,w,ch]")
        h, w = int(img.shape[0]), int(img.shape[1])
        ch = 1 if img.dim() == 2 else int(img.shape[2])
        return img.data_ptr(), w, h, ch, w * ch, img
    raise TypeError(f"unsupported image type {type(img)!r}")


def _buf_ptr(o) -> int:
    """Address of an output buffer: numpy (host) or a torch tensor (host or CUDA)."""
    return o.data_ptr() if hasattr(o, "data_ptr") else o.ctypes.data


def _buf_size(o) -> int:
    return int(o.numel()) if hasattr(o, "numel") else int(o.size)


def _dptr(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _result(cap: int):
    buf = np.zeros(max(1, cap), np.float64)
    r = OmrResultC(0.0, 0, 0, _dptr(buf), len(buf))
    return r, buf


def _unpack(r, buf) -> OmrResult:
    n = min(r.n_candidates, len(buf))
    return OmrResult(float(r.angle), int(r.status), [float(x) for x in buf[:n]])


# --------------------------------------------------------------------------- pure host helpers
def rotation_matrix_2d(cx, cy, angle, scale) -> np.ndarray:
    M = np.zeros(6, np.float64)
    _lib.load().oics_rotation_matrix_2d(C.c_float(cx), C.c_float(cy), angle, scale, _dptr(M))
    return M


def angle_half_range(max_angle: int, step: float) -> int:
    return int(_lib.load().oics_angle_half_range(int(max_angle), float(step)))


def sweep_matrices(w, h, max_angle, step, scale=1.0):
    """-> (N, angles, M[2N, 6])"""
    N = angle_half_range(max_angle, step)
    M = np.zeros((max(1, 2 * N), 6), np.float64)
    n = C.c_int32()
    check(_lib.load().oics_sweep_matrices(w, h, int(max_angle), float(step), float(scale), _dptr(M), 2 * N, C.byref(n)))
    return N, [float(d) * step for d in range(-N, N)], M[:2 * N]


def rotate_geometry(cols, rows, angle, scale=1.0, clip=CLIP_DEFAULT):
    M = np.zeros(6, np.float64)
    dw, dh = C.c_int32(), C.c_int32()
    check(_lib.load().oics_rotate_geometry(cols, rows, float(angle), float(scale), clip, _dptr(M), C.byref(dw), C.byref(dh)))
    return M, dw.value, dh.value


def shard_layout(n_angles, world) -> int:
    per = C.c_int32()
    check(_lib.load().oics_shard_layout(int(n_angles), int(world), C.byref(per)))
    return per.value


def projection_standard_deviations_static(d_bin, d_sd, stream=0) -> None:
    """oics_projection_standard_deviations_static (no counterpart in the reference): the std-devs of a
    DEVICE-resident binary image into a DEVICE double[2] (v_sd, h_sd), without any allocation and without
    a context.  `d_bin`, `d_sd`: objects with data_ptr() (torch CUDA tensors) on the current device;
    at most 16384 x 16384; enqueued on `stream` (a raw cudaStream_t), not synchronised."""
    h, w = int(d_bin.shape[0]), int(d_bin.shape[1])
    check(_lib.load().oics_projection_standard_deviations_static(
        C.c_void_p(d_bin.data_ptr()), w, h, int(d_bin.stride(0)), C.c_void_p(d_sd.data_ptr()), C.c_void_p(stream)))


def shard_unpack(sd, n_angles, world):
    """Gathered exchange buffer sd[world][2][per] -> (v_sd[n], h_sd[n])."""
    sd = np.ascontiguousarray(sd, np.float64)
    v, h = np.zeros(n_angles, np.float64), np.zeros(n_angles, np.float64)
    check(_lib.load().oics_shard_unpack(_dptr(sd), int(n_angles), int(world), _dptr(v), _dptr(h)))
    return v, h


def projection_scale(w, h, max_w, max_h) -> float:
    return float(_lib.load().oics_projection_scale(w, h, max_w, max_h))


def rule_a(h_sd, v_sd, N, step) -> OmrResult:
    h_sd = np.ascontiguousarray(h_sd, np.float64)
    v_sd = np.ascontiguousarray(v_sd, np.float64)
    r, buf = _result(2 * N)
    check(_lib.load().oics_rule_a(_dptr(h_sd), _dptr(v_sd), N, float(step), C.byref(r)))
    return _unpack(r, buf)


def rule_a_batch(h_sd, v_sd, N, step):
    """h_sd, v_sd: [n_cards, 2N] -> (angles, statuses, n_candidates) arrays."""
    h_sd = np.ascontiguousarray(h_sd, np.float64)
    v_sd = np.ascontiguousarray(v_sd, np.float64)
    n = h_sd.shape[0] if h_sd.ndim == 2 else 1
    angles = np.zeros(n, np.float64)
    stats = np.zeros(n, np.int32)
    ncand = np.zeros(n, np.int32)
    ip = C.POINTER(C.c_int32)
    check(_lib.load().oics_rule_a_batch(_dptr(h_sd), _dptr(v_sd), n, N, float(step), _dptr(angles),
                                        stats.ctypes.data_as(ip), ncand.ctypes.data_as(ip)))
    return angles, stats, ncand


def rule_b(v_sd, h_sd, N, step) -> float:
    v_sd = np.ascontiguousarray(v_sd, np.float64)
    h_sd = np.ascontiguousarray(h_sd, np.float64)
    a = C.c_double()
    check(_lib.load().oics_rule_b(_dptr(v_sd), _dptr(h_sd), N, float(step), C.byref(a)))
    return a.value


def lines_angle(lines, variant) -> OmrResult:
    lines = np.ascontiguousarray(lines, np.int32).reshape(-1, 4)
    r, buf = _result(len(lines))
    check(_lib.load().oics_lines_angle(lines.ctypes.data_as(C.POINTER(C.c_int32)), len(lines), variant, C.byref(r)))
    return _unpack(r, buf)


def reconcile(p: OmrResult, edges_angle: float):
    buf = np.ascontiguousarray(p.candidates if p.candidates else [0.0], np.float64)
    r = OmrResultC(p.angle, p.status, len(p.candidates), _dptr(buf), len(p.candidates))
    a, need = C.c_double(), C.c_int32()
    check(_lib.load().oics_reconcile(C.byref(r), float(edges_angle), C.byref(a), C.byref(need)))
    return a.value, bool(need.value)


def hough_dims(w, h, rho=1.0, theta=math.pi / 180.0):
    na, nr = C.c_int32(), C.c_int32()
    check(_lib.load().oics_hough_dims(w, h, rho, theta, C.byref(na), C.byref(nr)))
    return na.value, nr.value


class PinnedArray:
    """A page-locked uint8 host buffer (oics_alloc_pinned) exposed as a numpy array: `.array`.
    Keep the object alive as long as views of the array are in use."""

    def __init__(self, nbytes: int):
        self._lib = _lib.load()
        self._p = C.c_void_p()
        check(self._lib.oics_alloc_pinned(int(nbytes), C.byref(self._p)))
        self.array = np.ctypeslib.as_array(C.cast(self._p, C.POINTER(C.c_uint8)), shape=(int(nbytes),))

    def close(self):
        if getattr(self, "_p", None) is not None and self._p.value:
            self.array = None
            self._lib.oics_free_pinned(self._p)
            self._p = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


# --------------------------------------------------------------------------- device context
class Context:
    """One CUDA context handle (oics_ctx): device scratch + stream.  Use one per host thread."""

    def __init__(self, device: int = 0):
        self._lib = _lib.load()
        self._h = C.c_void_p()
        check(self._lib.oics_create(device, C.byref(self._h)))
        self.device = device

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            self._lib.oics_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # -- plumbing
    def set_stream(self, stream_ptr: int | None):
        check(self._lib.oics_set_stream(self._h, C.c_void_p(stream_ptr or 0)))

    def synchronize(self):
        check(self._lib.oics_synchronize(self._h))

    @property
    def launch_count(self) -> int:
        return int(self._lib.oics_launch_count(self._h))

    def profile(self, enable=True):
        """Opt-in per-kernel CUDA-event timing of this context's launches (oics_profile)."""
        check(self._lib.oics_profile(self._h, 1 if enable else 0))

    def profile_report(self):
        """-> list of {"kernel", "launches", "ms", "algorithmic_bytes"} since profile(True)."""
        import json
        need = C.c_size_t()
        cap = 1 << 16
        while True:
            buf = C.create_string_buffer(cap)
            rc = self._lib.oics_profile_report(self._h, buf, cap, C.byref(need))
            if rc == _lib.ERR_CAPACITY:
                cap = need.value + 64
                continue
            check(rc)
            return json.loads(buf.value.decode())

    def last_sweep_kernel_ms(self):
        ms, n = C.c_double(), C.c_int32()
        check(self._lib.oics_last_sweep_kernel_ms(self._h, C.byref(ms), C.byref(n)))
        return ms.value, n.value

    def probe_l2_read(self, nbytes=32 << 20, reps=64) -> float:
        """GB/s of L2-resident 16-byte reads over the whole device (bench roofline denominator)."""
        g = C.c_double()
        check(self._lib.oics_probe_l2_read(self._h, C.c_size_t(nbytes), reps, C.byref(g)))
        return g.value

    def probe_smem_read(self, iters=4096) -> float:
        """GB/s of shared-memory 16-byte reads over the whole device (bench roofline denominator)."""
        g = C.c_double()
        check(self._lib.oics_probe_smem_read(self._h, iters, C.byref(g)))
        return g.value

    # -- P0
    def preprocess(self, bgr, mode=PRE_OMR, max_w=0, max_h=0, resize_scale=1.0):
        """-> (binary [oh, ow] uint8, scale)"""
        p, w, h, ch, stride, keep = _describe(bgr)
        if ch != 3:
            raise ValueError("pre-processing expects a 3-channel image")
        scale = projection_scale(w, h, max_w, max_h) if mode == PRE_OMR else float(resize_scale)
        ow = w if scale == 1.0 else int(float(w) * scale)
        oh = h if scale == 1.0 else int(float(h) * scale)
        out = np.empty((max(oh, 1), max(ow, 1)), np.uint8)
        rw, rh, rs = C.c_int32(), C.c_int32(), C.c_double()
        check(self._lib.oics_preprocess(self._h, p, w, h, stride, mode, max_w, max_h, float(resize_scale),
                                        out.ctypes.data, out.size, C.byref(rw), C.byref(rh), C.byref(rs)))
        return out[:rh.value, :rw.value], rs.value

    # -- P2-P4
    def projection_sweep(self, binary, M, want_sums=False):
        """-> (v_sd, h_sd[, sums])"""
        p, w, h, ch, stride, keep = _describe(binary)
        M = np.ascontiguousarray(M, np.float64).reshape(-1, 6)
        n = len(M)
        v, hh = np.zeros(n, np.float64), np.zeros(n, np.float64)
        sums = np.zeros((n, 4), np.int64) if want_sums else None
        check(self._lib.oics_projection_sweep(self._h, p, w, h, stride, _dptr(M), n, v.ctypes.data, hh.ctypes.data,
                                              sums.ctypes.data if want_sums else None))
        return (v, hh, sums) if want_sums else (v, hh)

    def projection_sweep_batch(self, bins, n_cards, w, h, M, card_stride=None, stride=None, out=None):
        """bins: device pointer (int) or numpy [n, h, w]; -> (v_sd, h_sd) as [n_cards, n_angles]."""
        M = np.ascontiguousarray(M, np.float64).reshape(-1, 6)
        n = len(M)
        if isinstance(bins, np.ndarray):
            bins = np.ascontiguousarray(bins, np.uint8)
            ptr, keep = bins.ctypes.data, bins
        elif hasattr(bins, "data_ptr"):
            ptr, keep = bins.data_ptr(), bins
        else:
            ptr, keep = int(bins), None
        stride = stride or w
        card_stride = card_stride or stride * h
        if out is None:
            v, hh = np.zeros((n_cards, n), np.float64), np.zeros((n_cards, n), np.float64)
        else:
            v, hh = out
        check(self._lib.oics_projection_sweep_batch(self._h, ptr, n_cards, card_stride, w, h, stride, _dptr(M), n,
                                                    v.ctypes.data, hh.ctypes.data))
        return v, hh

    def sweep_sharded(self, binary, M, rank, world, sd=None):
        """oics_sweep_sharded (config 5): evaluates this rank's slice of M [n, 6] into its block of the
        exchange buffer sd[world][2][per] (numpy float64, or a CUDA tensor for an in-place NCCL all-gather).
        -> (sd, first, count, per)"""
        p, w, h, ch, stride, keep = _describe(binary)
        M = np.ascontiguousarray(M, np.float64).reshape(-1, 6)
        n = len(M)
        per = shard_layout(n, world)
        if sd is None:
            sd = np.zeros((world, 2, per), np.float64)
        ptr = sd.data_ptr() if hasattr(sd, "data_ptr") else sd.ctypes.data
        first, count = C.c_int32(), C.c_int32()
        check(self._lib.oics_sweep_sharded(self._h, p, w, h, stride, _dptr(M), n, int(rank), int(world), C.c_void_p(ptr),
                                           C.byref(first), C.byref(count)))
        return sd, first.value, count.value, per

    def projection_counts(self, binary, M):
        p, w, h, ch, stride, keep = _describe(binary)
        M = np.ascontiguousarray(M, np.float64).reshape(6)
        row, col = np.zeros(h, np.int32), np.zeros(w, np.int32)
        check(self._lib.oics_projection_counts(self._h, p, w, h, stride, _dptr(M), row.ctypes.data, col.ctypes.data))
        return row, col

    def projection_standard_deviations(self, binary):
        """transfer::get_projection_standard_deviations -> (vertical_sd, horizontal_sd)."""
        p, w, h, ch, stride, keep = _describe(binary)
        v, hh = C.c_double(), C.c_double()
        check(self._lib.oics_projection_standard_deviations(self._h, p, w, h, stride, C.byref(v), C.byref(hh)))
        return v.value, hh.value

    # -- estimators
    def result_from_projection(self, bgr, max_angle, step, max_w, max_h) -> OmrResult:
        p, w, h, ch, stride, keep = _describe(bgr)
        r, buf = _result(2 * angle_half_range(max_angle, step))
        check(self._lib.oics_result_from_projection(self._h, p, w, h, stride, int(max_angle), float(step), max_w, max_h,
                                                    C.byref(r)))
        return _unpack(r, buf)

    def angle_with_projections(self, bgr, max_angle, step, resize_scale, threads=1) -> float:
        p, w, h, ch, stride, keep = _describe(bgr)
        a = C.c_double()
        check(self._lib.oics_angle_with_projections(self._h, p, w, h, stride, int(max_angle), float(step),
                                                    float(resize_scale), int(threads), C.byref(a)))
        return a.value

    def result_from_edges_detection(self, img, min_len, max_gap) -> OmrResult:
        p, w, h, ch, stride, keep = _describe(img)
        r, buf = _result(1 << 16)
        check(self._lib.oics_result_from_edges_detection(self._h, p, ch, w, h, stride, float(min_len), float(max_gap),
                                                         C.byref(r)))
        return _unpack(r, buf)

    def result_from_edges_detection_batch(self, imgs, min_len, max_gap):
        """omr::get_result_from_edges_detection for n equally sized cards in flight together
        (numpy [n, h, w(, 3)] or a CUDA tensor of that shape) -> (angles, result statuses, error statuses)."""
        if isinstance(imgs, np.ndarray):
            imgs = np.ascontiguousarray(imgs, np.uint8)
            ptr = imgs.ctypes.data
        else:
            ptr = imgs.data_ptr()
        n, h, w = int(imgs.shape[0]), int(imgs.shape[1]), int(imgs.shape[2])
        ch = 1 if len(imgs.shape) == 3 else int(imgs.shape[3])
        angles = np.zeros(n, np.float64)
        rstat = np.zeros(n, np.int32)
        stat = np.zeros(n, np.int32)
        ip = C.POINTER(C.c_int32)
        check(self._lib.oics_result_from_edges_detection_batch(self._h, ptr, n, w * h * ch, ch, w, h, w * ch, float(min_len),
                                                               float(max_gap), _dptr(angles), rstat.ctypes.data_as(ip),
                                                               stat.ctypes.data_as(ip)))
        return angles, rstat, stat

    def angle_with_hough(self, gray, min_len, max_gap) -> float:
        p, w, h, ch, stride, keep = _describe(gray)
        a = C.c_double()
        check(self._lib.oics_angle_with_hough(self._h, p, w, h, stride, float(min_len), float(max_gap), C.byref(a)))
        return a.value

    def angle_with_fft(self, gray, canny1, canny2, min_len, max_gap) -> float:
        p, w, h, ch, stride, keep = _describe(gray)
        a = C.c_double()
        check(self._lib.oics_angle_with_fft(self._h, p, w, h, stride, float(canny1), float(canny2), float(min_len),
                                            float(max_gap), C.byref(a)))
        return a.value

    def result_from_fourier_transform(self, bgr, canny_weak, canny_strong, min_len, max_gap) -> OmrResult:
        p, w, h, ch, stride, keep = _describe(bgr)
        r, buf = _result(1 << 16)
        check(self._lib.oics_result_from_fourier_transform(self._h, p, w, h, stride, float(canny_weak),
                                                           float(canny_strong), float(min_len), float(max_gap),
                                                           C.byref(r)))
        return _unpack(r, buf)

    # -- building blocks
    def canny(self, img, lo, hi) -> np.ndarray:
        p, w, h, ch, stride, keep = _describe(img)
        out = np.empty((h, w), np.uint8)
        check(self._lib.oics_canny(self._h, p, ch, w, h, stride, float(lo), float(hi), out.ctypes.data, w))
        return out

    def hough_accumulate(self, edges, rho=1.0, theta=math.pi / 180.0) -> np.ndarray:
        p, w, h, ch, stride, keep = _describe(edges)
        na, nr = hough_dims(w, h, rho, theta)
        acc = np.zeros((na, nr), np.int32)
        check(self._lib.oics_hough_accumulate(self._h, p, w, h, stride, rho, theta, acc.ctypes.data))
        return acc

    def hough_peaks(self, accum, threshold) -> np.ndarray:
        """Local maxima of a standard accumulator [numangle, numrho] in cv::HoughLines' order
        -> [n, 3] int32 rows {theta index, rho index, count}."""
        if hasattr(accum, "data_ptr"):
            ptr, (na, nr) = accum.data_ptr(), accum.shape
        else:
            accum = np.ascontiguousarray(accum, np.int32)
            ptr, (na, nr) = accum.ctypes.data, accum.shape
        cap = 1 << 12
        while True:
            peaks = np.zeros((cap, 3), np.int32)
            n = C.c_int32()
            rc = self._lib.oics_hough_peaks(self._h, C.c_void_p(ptr), int(na), int(nr), int(threshold), peaks.ctypes.data,
                                            cap, C.byref(n))
            if rc == _lib.ERR_CAPACITY:
                cap = n.value
                continue
            check(rc)
            return peaks[:n.value].copy()

    def hough_lines_p(self, edges, threshold, min_len, max_gap, rho=1.0, theta=math.pi / 180.0) -> np.ndarray:
        p, w, h, ch, stride, keep = _describe(edges)
        cap = 1 << 14
        while True:
            lines = np.zeros((cap, 4), np.int32)
            n = C.c_int32()
            rc = self._lib.oics_hough_lines_p(self._h, p, w, h, stride, rho, theta, int(threshold), float(min_len),
                                              float(max_gap), lines.ctypes.data_as(C.POINTER(C.c_int32)), cap, C.byref(n))
            if rc == _lib.ERR_CAPACITY:
                cap = n.value
                continue
            check(rc)
            return lines[:n.value].copy()

    def fft_image(self, gray, want_magnitude=False):
        p, w, h, ch, stride, keep = _describe(gray)
        log8 = np.empty((h, w), np.uint8)
        mag8 = np.empty((h, w), np.uint8) if want_magnitude else None
        check(self._lib.oics_fft_image(self._h, p, w, h, stride, mag8.ctypes.data if want_magnitude else None,
                                       log8.ctypes.data))
        return (mag8, log8)

    # -- W1
    def warp_affine(self, src, M, dsize, flags=INTER_NEAREST, border=WHITE, out=None):
        p, w, h, ch, stride, keep = _describe(src)
        dw, dh = dsize
        M = np.ascontiguousarray(M, np.float64).reshape(6)
        b = (C.c_uint8 * 4)(*[int(v) & 255 for v in border])
        if out is None:
            out = np.empty((dh, dw) if ch == 1 else (dh, dw, ch), np.uint8)
        op, ow, oh, och, ostride, okeep = _describe(out)
        check(self._lib.oics_warp_affine(self._h, p, ch, w, h, stride, _dptr(M), flags, b, op, dw, dh, ostride))
        return out

    def rotate_mat(self, src, angle, scale=1.0, flags=INTER_NEAREST, border=WHITE, clip=CLIP_DEFAULT):
        """transfer::rotate_mat (transfer.rs:459-467)."""
        p, w, h, ch, stride, keep = _describe(src)
        _, dw, dh = rotate_geometry(w, h, angle, scale, clip)
        out = np.empty((dh, dw) if ch == 1 else (dh, dw, ch), np.uint8)
        b = (C.c_uint8 * 4)(*[int(v) & 255 for v in border])
        rw, rh = C.c_int32(), C.c_int32()
        check(self._lib.oics_rotate(self._h, p, ch, w, h, stride, float(angle), float(scale), flags, b, clip,
                                    out.ctypes.data, out.size, C.byref(rw), C.byref(rh)))
        return out

    # -- T1
    def scale(self, src, scale):
        """TransformableMatrix::scale_self (transfer.rs:66-91) -> scaled image."""
        p, w, h, ch, stride, keep = _describe(src)
        ow = w if scale == 1.0 else int(float(w) * scale)
        oh = h if scale == 1.0 else int(float(h) * scale)
        out = np.empty((max(oh, 1), max(ow, 1)) if ch == 1 else (max(oh, 1), max(ow, 1), ch), np.uint8)
        rw, rh = C.c_int32(), C.c_int32()
        check(self._lib.oics_scale(self._h, p, ch, w, h, stride, float(scale), out.ctypes.data, out.size, C.byref(rw),
                                   C.byref(rh)))
        return out

    def shrink_to(self, src, max_w, max_h):
        """TransformableMatrix::shrink_to (transfer.rs:93-126) -> (image, scale used); never upscales."""
        p, w, h, ch, stride, keep = _describe(src)
        s = min(1.0, projection_scale(w, h, max_w, max_h))
        ow = w if s == 1.0 else int(float(w) * s)
        oh = h if s == 1.0 else int(float(h) * s)
        out = np.empty((max(oh, 1), max(ow, 1)) if ch == 1 else (max(oh, 1), max(ow, 1), ch), np.uint8)
        rw, rh, rs = C.c_int32(), C.c_int32(), C.c_double()
        check(self._lib.oics_shrink_to(self._h, p, ch, w, h, stride, int(max_w), int(max_h), out.ctypes.data, out.size,
                                       C.byref(rw), C.byref(rh), C.byref(rs)))
        return out, rs.value

    # -- N2: imwrite
    def jpeg_encode(self, bgr, quality=100, out=None) -> bytes:
        """TransformableMatrix::im_write(path, JPEG, quality) (transfer.rs:162-179) -> the file's bytes."""
        p, w, h, ch, stride, keep = _describe(bgr)
        if ch != 3:
            raise ValueError("jpeg_encode expects a BGR image")
        if out is None:
            out = np.empty(4096 + 4 * w * h, np.uint8)
        n = C.c_size_t()
        check(self._lib.oics_jpeg_encode(self._h, p, w, h, stride, int(quality), out.ctypes.data, out.size, C.byref(n)))
        return out[:n.value].tobytes()

    def jpeg_info(self, data):
        """Header of a JPEG file's bytes -> (width, height, components); host only."""
        buf = np.frombuffer(data, np.uint8) if not isinstance(data, np.ndarray) else data
        w, h, nc = C.c_int32(), C.c_int32(), C.c_int32()
        check(self._lib.oics_jpeg_info(buf.ctypes.data, buf.size, C.byref(w), C.byref(h), C.byref(nc)))
        return w.value, h.value, nc.value

    def jpeg_decode(self, data, out=None):
        """imgcodecs::imread(path, IMREAD_COLOR) (omr.rs:349, transfer.rs:43) of a baseline JPEG file's
        bytes -> BGR image, decoded on the GPU.  `out`: optional flat uint8 buffer (e.g. pinned memory,
        or a torch CUDA tensor: the pixels then stay on the device)."""
        buf = np.frombuffer(data, np.uint8) if not isinstance(data, np.ndarray) else data
        w, h, nc = C.c_int32(), C.c_int32(), C.c_int32()
        check(self._lib.oics_jpeg_info(buf.ctypes.data, buf.size, C.byref(w), C.byref(h), C.byref(nc)))
        need = w.value * h.value * 3
        if out is None:
            out = np.empty(need, np.uint8)
        if hasattr(out, "data_ptr"):
            ptr, cap = out.data_ptr(), out.numel()
        else:
            ptr, cap = out.ctypes.data, out.size
        check(self._lib.oics_jpeg_decode(self._h, buf.ctypes.data, buf.size, ptr, 0, cap, C.byref(w), C.byref(h)))
        if hasattr(out, "data_ptr"):
            return out[:need].view(h.value, w.value, 3)
        return out[:need].reshape(h.value, w.value, 3)

    # -- pipeline
    def correct_default(self, bgr, max_angle, step, max_w, max_h, hough_min_len, hough_max_gap, want_image=True,
                        out=None):
        """omr::correct_default on memory buffers -> (angle, need_check, deskewed BGR or None).
        `out`: optional flat uint8 buffer for the deskewed card (numpy, e.g. a view of pinned memory,
        or anything api._describe-like with .ctypes.data/.size); the returned image is a view of it.
        A caller that deskews card after card should pass one: a fresh pageable 30 MB array costs
        more in page faults than the whole GPU path."""
        p, w, h, ch, stride, keep = _describe(bgr)
        cap = 0
        if want_image:
            # CONTAIN output is at most (w+h) x (w+h)
            side = int(math.ceil(math.hypot(w, h))) + 2
            if out is None:
                out = np.empty(side * side * 3, np.uint8)
            cap = out.size
        dw, dh, a, need = C.c_int32(), C.c_int32(), C.c_double(), C.c_int32()
        check(self._lib.oics_correct_default(self._h, p, w, h, stride, int(max_angle), float(step), max_w, max_h,
                                             float(hough_min_len), float(hough_max_gap),
                                             out.ctypes.data if want_image else None, cap, C.byref(dw), C.byref(dh),
                                             C.byref(a), C.byref(need)))
        img = out[:dw.value * dh.value * 3].reshape(dh.value, dw.value, 3) if want_image else None
        return a.value, bool(need.value), img

    def correct_batch(self, bgrs, max_angle, step, max_w, max_h, hough_min_len, hough_max_gap, outs=None,
                      jpeg_quality=None):
        """bgrs: numpy [n, h, w, 3] (host) or CUDA tensor of that shape.
        outs: optional list of n writable uint8 numpy buffers for the deskewed cards.
        jpeg_quality: None -> the buffers receive BGR pixels; 1..100 -> the JPEG file imwrite would write
        (omr.rs:447-448 uses 100) and a fifth return value holds the file sizes.
        -> (angles, need_checks, statuses, dst sizes [n, 2][, file sizes])"""
        if isinstance(bgrs, np.ndarray):
            bgrs = np.ascontiguousarray(bgrs, np.uint8)
            ptr = bgrs.ctypes.data
        else:
            ptr = bgrs.data_ptr()
        n, h, w = int(bgrs.shape[0]), int(bgrs.shape[1]), int(bgrs.shape[2])
        angles = np.zeros(n, np.float64)
        needs = np.zeros(n, np.int32)
        stats = np.zeros(n, np.int32)
        sizes = np.zeros((2, n), np.int32)
        dsts = caps = None
        if outs is not None:
            dsts = (C.c_void_p * n)(*[_buf_ptr(o) for o in outs])
            caps = (C.c_size_t * n)(*[_buf_size(o) for o in outs])
        ip = C.POINTER(C.c_int32)
        if jpeg_quality is None:
            check(self._lib.oics_correct_batch(self._h, ptr, n, w * h * 3, w, h, w * 3, int(max_angle), float(step), max_w,
                                               max_h, float(hough_min_len), float(hough_max_gap), dsts, caps,
                                               sizes[0].ctypes.data_as(ip), sizes[1].ctypes.data_as(ip), _dptr(angles),
                                               needs.ctypes.data_as(ip), stats.ctypes.data_as(ip)))
            return angles, needs.astype(bool), stats, sizes.T.copy()
        fsz = (C.c_size_t * n)()
        check(self._lib.oics_correct_batch_jpeg(self._h, ptr, n, w * h * 3, w, h, w * 3, int(max_angle), float(step), max_w,
                                                max_h, float(hough_min_len), float(hough_max_gap), int(jpeg_quality), dsts,
                                                caps, fsz, sizes[0].ctypes.data_as(ip), sizes[1].ctypes.data_as(ip),
                                                _dptr(angles), needs.ctypes.data_as(ip), stats.ctypes.data_as(ip)))
        return angles, needs.astype(bool), stats, sizes.T.copy(), np.array(list(fsz), np.int64)

    def correct_batch_files(self, files, max_angle, step, max_w, max_h, hough_min_len, hough_max_gap, outs=None,
                            jpeg_quality=100):
        """omr::correct_default from file bytes to file bytes: `files` is a list of uint8 numpy arrays (JPEG
        files, ideally views of pinned memory), all cards of the first file's size.  jpeg_quality 0 -> the
        buffers in `outs` receive BGR pixels, 1..100 -> the JPEG file imwrite would write.
        -> (angles, need_checks, statuses, dst sizes [n, 2], file sizes)"""
        n = len(files)
        files = [f if isinstance(f, np.ndarray) else np.frombuffer(f, np.uint8) for f in files]
        fptr = (C.c_void_p * n)(*[f.ctypes.data for f in files])
        flen = (C.c_size_t * n)(*[f.size for f in files])
        angles = np.zeros(n, np.float64)
        needs = np.zeros(n, np.int32)
        stats = np.zeros(n, np.int32)
        sizes = np.zeros((2, n), np.int32)
        dsts = caps = None
        if outs is not None:
            dsts = (C.c_void_p * n)(*[_buf_ptr(o) for o in outs])
            caps = (C.c_size_t * n)(*[_buf_size(o) for o in outs])
        ip = C.POINTER(C.c_int32)
        fsz = (C.c_size_t * n)()
        check(self._lib.oics_correct_batch_files(self._h, fptr, flen, n, int(max_angle), float(step), max_w, max_h,
                                                 float(hough_min_len), float(hough_max_gap), int(jpeg_quality), dsts, caps, fsz,
                                                 sizes[0].ctypes.data_as(ip), sizes[1].ctypes.data_as(ip), _dptr(angles),
                                                 needs.ctypes.data_as(ip), stats.ctypes.data_as(ip)))
        return angles, needs.astype(bool), stats, sizes.T.copy(), np.array(list(fsz), np.int64)


_default = {}
_default_lock = threading.Lock()


def default_context(device: int = 0) -> Context:
    """A lazily created per-(thread, device) context for the module-level mirror functions."""
    key = (threading.get_ident(), device)
    with _default_lock:
        ctx = _default.get(key)
        if ctx is None:
            ctx = _default[key] = Context(device)
        return ctx
