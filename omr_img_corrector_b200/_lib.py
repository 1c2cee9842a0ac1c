"""ctypes binding of liboics_cuda.so (the C ABI in include/oics_cuda.h).

Loading fails loudly: there is no Python or CPU fallback for any compute entry.
"""
from __future__ import annotations

import ctypes as C
import os

# The file-to-file batch driver decodes the cards of a sub-batch on one stream each (up to 32): with the CUDA
# default of 8 hardware work queues those streams alias and 7 decodes run at a time.  Only read at CUDA
# initialisation, so it is set when the package is imported (a value the caller exported wins); the library's
# own load-time constructor (context.cu) does the same for hosts that bind the C ABI directly.
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")

HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.path.join(HERE, "liboics_cuda.so")

OK = 0
ERR_INVALID_ARG, ERR_CUDA, ERR_NO_LINES, ERR_UNIMPLEMENTED, ERR_NOMEM, ERR_UNSUPPORTED, ERR_CAPACITY = range(-1, -8, -1)


class OicsError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"oics error {code}: {message}")
        self.code = code
        self.message = message


class OmrResultC(C.Structure):
    _fields_ = [("angle", C.c_double), ("status", C.c_int32), ("n_candidates", C.c_int32),
                ("candidates", C.POINTER(C.c_double)), ("cap", C.c_int32)]


_u8p = C.c_void_p            # image pointers: host numpy data or raw device addresses
_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int32)
_i32, _f64, _sz = C.c_int32, C.c_double, C.c_size_t
_ctx = C.c_void_p

# name -> (restype, argtypes); every symbol the header declares
SIGNATURES = {
    "oics_create": (_i32, [_i32, C.POINTER(_ctx)]),
    "oics_destroy": (None, [_ctx]),
    "oics_last_error": (C.c_char_p, []),
    "oics_version": (_i32, []),
    "oics_set_stream": (_i32, [_ctx, C.c_void_p]),
    "oics_synchronize": (_i32, [_ctx]),
    "oics_launch_count": (C.c_int64, [_ctx]),
    "oics_alloc_pinned": (_i32, [_sz, C.POINTER(C.c_void_p)]),
    "oics_free_pinned": (None, [C.c_void_p]),
    "oics_profile": (_i32, [_ctx, _i32]),
    "oics_profile_report": (_i32, [_ctx, C.c_char_p, _sz, C.POINTER(_sz)]),
    "oics_last_sweep_kernel_ms": (_i32, [_ctx, _dp, _ip]),
    "oics_probe_l2_read": (_i32, [_ctx, _sz, _i32, _dp]),
    "oics_probe_smem_read": (_i32, [_ctx, _i32, _dp]),
    "oics_rotation_matrix_2d": (None, [C.c_float, C.c_float, _f64, _f64, _dp]),
    "oics_angle_half_range": (_i32, [C.c_uint32, _f64]),
    "oics_sweep_matrices": (_i32, [_i32, _i32, C.c_uint32, _f64, _f64, _dp, _i32, _ip]),
    "oics_shard_range": (_i32, [_i32, _i32, _i32, _ip, _ip]),
    "oics_shard_layout": (_i32, [_i32, _i32, _ip]),
    "oics_shard_unpack": (_i32, [_dp, _i32, _i32, _dp, _dp]),
    "oics_sweep_sharded": (_i32, [_ctx, _u8p, _i32, _i32, _i32, _dp, _i32, _i32, _i32, C.c_void_p, _ip, _ip]),
    "oics_hough_peaks": (_i32, [_ctx, C.c_void_p, _i32, _i32, _i32, C.c_void_p, _i32, _ip]),
    "oics_scale": (_i32, [_ctx, _u8p, _i32, _i32, _i32, _i32, _f64, _u8p, _sz, _ip, _ip]),
    "oics_shrink_to": (_i32, [_ctx, _u8p, _i32, _i32, _i32, _i32, _i32, _i32, _u8p, _sz, _ip, _ip, _dp]),
    "oics_rotate_geometry": (_i32, [_i32, _i32, _f64, _f64, _i32, _dp, _ip, _ip]),
    "oics_projection_scale": (_f64, [_i32, _i32, _i32, _i32]),
    "oics_preprocess": (_i32, [_ctx, _u8p, _i32, _i32, _i32, _i32, _i32, _i32, _f64, _u8p, _sz, _ip, _ip, _dp]),
    "oics_projection_sweep": (_i32, [_ctx, _u8p, _i32, _i32, _i32, _dp, _i32, C.c_void_p, C.c_void_p, C.c_void_p]),
    "oics_projection_sweep_batch": (_i32, [_ctx, _u8p, _i32, _sz, _i32, _i32, _i32, _dp, _i32, C.c_void_p, C.c_void_p]),
    "oics_projection_counts": (_i32, [_ctx, _u8p, _i32, _i32, _i32, _dp, C.c_void_p, C.c_void_p]),
    "oics_projection_standard_deviations": (_i32, [_ctx, _u8p, _i32, _i32, _i32, _dp, _dp]),
    "oics_projection_standard_deviations_static": (_i32, [_u8p, _i32, _i32, _i32, C.c_void_p, C.c_void_p]),
    "oics_rule_a": (_i32, [_dp, _dp, _i32, _f64, C.POINTER(OmrResultC)]),
    "oics_rule_a_batch": (_i32, [_dp, _dp, _i32, _i32, _f64, _dp, _ip, _ip]),
    "oics_rule_b": (_i32, [_dp, _dp, _i32, _f64, _dp]),
    "oics_result_from_projection": (_i32, [_ctx, _u8p, _i32, _i32, _i32, C.c_uint16, _f64, _i32, _i32, C.POINTER(OmrResultC)]),
    "oics_angle_with_projections": (_i32, [_ctx, _u8p, _i32, _i32, _i32, C.c_uint16, _f64, _f64, _sz, _dp]),
    "oics_result_from_edges_detection": (_i32, [_ctx, _u8p, _i32, _i32, _i32, _i32, _f64, _f64, C.POINTER(OmrResultC)]),
    "oics_result_from_edges_detection_batch": (_i32, [_ctx, _u8p, _i32, _sz, _i32, _i32, _i32, _i32, _f64, _f64, _dp, _ip, _ip]),
    "oics_angle_with_hough": (_i32, [_ctx, _u8p, _i32, _i32, _i32, _f64, _f64, _dp]),
    "oics_angle_with_fft": (_i32, [_ctx, _u8p, _i32, _i32, _i32, _f64, _f64, _f64, _f64, _dp]),
    "oics_result_from_fourier_transform": (_i32, [_ctx, _u8p, _i32, _i32, _i32, _f64, _f64, _f64, _f64, C.POINTER(OmrResultC)]),
    "oics_canny": (_i32, [_ctx, _u8p, _i32, _i32, _i32, _i32, _f64, _f64, _u8p, _i32]),
    "oics_hough_dims": (_i32, [_i32, _i32, _f64, _f64, _ip, _ip]),
    "oics_hough_accumulate": (_i32, [_ctx, _u8p, _i32, _i32, _i32, _f64, _f64, C.c_void_p]),
    "oics_hough_lines_p": (_i32, [_ctx, _u8p, _i32, _i32, _i32, _f64, _f64, _i32, _f64, _f64, _ip, _i32, _ip]),
    "oics_lines_angle": (_i32, [_ip, _i32, _i32, C.POINTER(OmrResultC)]),
    "oics_fft_image": (_i32, [_ctx, _u8p, _i32, _i32, _i32, _u8p, _u8p]),
    "oics_warp_affine": (_i32, [_ctx, _u8p, _i32, _i32, _i32, _i32, _dp, _i32, C.POINTER(C.c_uint8), _u8p, _i32, _i32, _i32]),
    "oics_rotate": (_i32, [_ctx, _u8p, _i32, _i32, _i32, _i32, _f64, _f64, _i32, C.POINTER(C.c_uint8), _i32, _u8p, _sz, _ip, _ip]),
    "oics_jpeg_encode": (_i32, [_ctx, _u8p, _i32, _i32, _i32, _i32, _u8p, _sz, C.POINTER(_sz)]),
    "oics_jpeg_info": (_i32, [_u8p, _sz, _ip, _ip, _ip]),
    "oics_jpeg_decode": (_i32, [_ctx, _u8p, _sz, _u8p, _sz, _sz, _ip, _ip]),
    "oics_reconcile": (_i32, [C.POINTER(OmrResultC), _f64, _dp, _ip]),
    "oics_correct_default": (_i32, [_ctx, _u8p, _i32, _i32, _i32, C.c_uint16, _f64, _i32, _i32, _f64, _f64, _u8p, _sz, _ip, _ip, _dp, _ip]),
    "oics_correct_batch": (_i32, [_ctx, _u8p, _i32, _sz, _i32, _i32, _i32, C.c_uint16, _f64, _i32, _i32, _f64, _f64,
                                  C.POINTER(C.c_void_p), C.POINTER(_sz), _ip, _ip, _dp, _ip, _ip]),
    "oics_correct_batch_files": (_i32, [_ctx, C.POINTER(C.c_void_p), C.POINTER(_sz), _i32, C.c_uint16, _f64, _i32, _i32, _f64, _f64, _i32,
                                        C.POINTER(C.c_void_p), C.POINTER(_sz), C.POINTER(_sz), _ip, _ip, _dp, _ip, _ip]),
    "oics_correct_batch_jpeg": (_i32, [_ctx, _u8p, _i32, _sz, _i32, _i32, _i32, C.c_uint16, _f64, _i32, _i32, _f64, _f64, _i32,
                                       C.POINTER(C.c_void_p), C.POINTER(_sz), C.POINTER(_sz), _ip, _ip, _dp, _ip, _ip]),
}

_lib = None


def load():
    """Returns the CDLL with argtypes set.  Raises if the library was not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(SO_PATH):
        raise ImportError(
            f"{SO_PATH} is missing: build it with `python -m omr_img_corrector_b200.build` "
            "(the product has no CPU fallback)")
    lib = C.CDLL(SO_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if the header and the .so disagree
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def last_error() -> str:
    return load().oics_last_error().decode("utf-8", "replace")


def check(rc: int) -> None:
    if rc != OK:
        raise OicsError(rc, last_error())
