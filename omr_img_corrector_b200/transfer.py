"""packages/lib/src/transfer.rs -- the hot subset: `TransformableMatrix` (the owning image
wrapper: load / from_matrix / get_mat / scale_self / shrink_to / get_bytes / im_write / clone),
rotate_mat, projection std-devs, gray / threshold helpers.  Every pixel operation is a call
into liboics_cuda.so; the numpy array held by the wrapper is the `cv::Mat`."""
from __future__ import annotations

import numpy as np

from . import api
from .types import ImageFormat, RotateClipStrategy


class TransformableMatrix:
    """transfer::TransformableMatrix (transfer.rs:16-277).  `scale_self` / `shrink_to` mutate and
    return self like the reference's `&mut Self` builders; `from_matrix` / `clone` deep-copy
    (transfer.rs:55-59, 181-185)."""

    def __init__(self, matrix: np.ndarray | None = None, ctx: "api.Context | None" = None):
        self.matrix = matrix if matrix is not None else np.zeros((0, 0), np.uint8)   # default(): empty Mat
        self._ctx = ctx

    # -- construction (transfer.rs:30-64)
    @classmethod
    def default(cls) -> "TransformableMatrix":
        return cls()

    @classmethod
    def from_matrix(cls, mat: np.ndarray, ctx=None) -> "TransformableMatrix":
        return cls(np.array(mat, np.uint8, copy=True), ctx)

    @classmethod
    def new(cls, filename: str, flags: int = 1, ctx=None) -> "TransformableMatrix":
        """TransformableMatrix::new(filename, flags) = imread (transfer.rs:61-64); IMREAD_COLOR (1) only,
        baseline JPEG files (the device decoder's subset)."""
        return cls(ctx=ctx).load_mat(filename, flags)

    def load_mat(self, filename: str, flags: int = 1) -> "TransformableMatrix":
        if flags != 1:
            raise api.OicsError(-6, "only IMREAD_COLOR (1) is supported")
        data = np.fromfile(filename, np.uint8)
        self.matrix = np.array(self._context().jpeg_decode(data), copy=True)
        return self

    def _context(self) -> "api.Context":
        return self._ctx or api.default_context()

    # -- accessors
    def get_mat(self) -> np.ndarray:
        return self.matrix

    def get_bytes(self) -> bytes:
        """data_bytes() (transfer.rs:156-158)."""
        return np.ascontiguousarray(self.matrix).tobytes()

    def clone(self) -> "TransformableMatrix":
        return TransformableMatrix(self.matrix.copy(), self._ctx)

    to_owned = clone

    # -- the hot members
    def scale_self(self, scale: float) -> "TransformableMatrix":
        """transfer.rs:66-91: INTER_AREA below 1, unchanged at 1; enlarging is outside the accelerated path."""
        if scale == 1.0:
            return self
        self.matrix = self._context().scale(self.matrix, float(scale))
        return self

    def shrink_to(self, max_width: int, max_height: int) -> "TransformableMatrix":
        """transfer.rs:93-126: never upscales."""
        out, s = self._context().shrink_to(self.matrix, int(max_width), int(max_height))
        if s < 1.0:
            self.matrix = out
        return self

    def im_write(self, filename: str, format: ImageFormat = ImageFormat.JPEG, quality: int = 100) -> bool:
        """transfer.rs:162-179; JPEG only (the device encoder), BGR images."""
        if format != ImageFormat.JPEG:
            raise api.OicsError(-6, "only ImageFormat.JPEG is written on the device")
        data = self._context().jpeg_encode(self.matrix, int(quality))
        with open(filename, "wb") as fh:
            fh.write(data)
        return True


def rotate_mat(src, angle, scale, flags, border_mode=0, border_value=api.WHITE,
               clip_strategy=RotateClipStrategy.DEFAULT, ctx=None):
    """transfer::rotate_mat (transfer.rs:459-467).  border_mode must be BORDER_CONSTANT (0),
    the only mode the reference passes.  `src` may be a TransformableMatrix (-> TransformableMatrix)
    or a bare image (-> image)."""
    if border_mode != 0:
        raise api.OicsError(-6, "only BORDER_CONSTANT is supported")
    wrapped = isinstance(src, TransformableMatrix)
    ctx = ctx or (src._ctx if wrapped else None) or api.default_context()
    out = ctx.rotate_mat(src.matrix if wrapped else src, angle, scale, flags, border_value, int(clip_strategy))
    return TransformableMatrix(out, ctx) if wrapped else out


def _mat(src):
    return src.matrix if isinstance(src, TransformableMatrix) else src


def get_projection_standard_deviations(src, ctx=None):
    """transfer::get_projection_standard_deviations (transfer.rs:527-536)
    -> (vertical_sd, horizontal_sd)."""
    ctx = ctx or api.default_context()
    return ctx.projection_standard_deviations(_mat(src))


def transfer_rgb_image_to_thresh_binary(src, ctx=None):
    """transfer_rgb_image_to_gray_image + transfer_gray_image_to_thresh_binary
    (transfer.rs:283-301) fused: gray then threshold 127."""
    ctx = ctx or api.default_context()
    return ctx.preprocess(_mat(src), api.PRE_PROJECTION, resize_scale=1.0)[0]
