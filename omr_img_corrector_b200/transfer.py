"""packages/lib/src/transfer.rs -- the hot subset: rotate_mat, projection std-devs,
gray / threshold helpers (via the pre-processing kernel)."""
from . import api
from .types import RotateClipStrategy


def rotate_mat(src, angle, scale, flags, border_mode=0, border_value=api.WHITE,
               clip_strategy=RotateClipStrategy.DEFAULT, ctx=None):
    """transfer::rotate_mat (transfer.rs:459-467).  border_mode must be BORDER_CONSTANT (0),
    the only mode the reference passes."""
    if border_mode != 0:
        raise api.OicsError(-6, "only BORDER_CONSTANT is supported")
    ctx = ctx or api.default_context()
    return ctx.rotate_mat(src, angle, scale, flags, border_value, int(clip_strategy))


def get_projection_standard_deviations(src, ctx=None):
    """transfer::get_projection_standard_deviations (transfer.rs:527-536)
    -> (vertical_sd, horizontal_sd)."""
    ctx = ctx or api.default_context()
    return ctx.projection_standard_deviations(src)


def transfer_rgb_image_to_thresh_binary(src, ctx=None):
    """transfer_rgb_image_to_gray_image + transfer_gray_image_to_thresh_binary
    (transfer.rs:283-301) fused: gray then threshold 127."""
    ctx = ctx or api.default_context()
    return ctx.preprocess(src, api.PRE_PROJECTION, resize_scale=1.0)[0]
