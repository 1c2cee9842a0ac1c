"""packages/lib/src/projection.rs."""
from . import api


def get_angle_with_projections(src_img, max_angle, step, resize_scale, threads=1, ctx=None):
    """projection::get_angle_with_projections (projection.rs:17-23).  `threads` is accepted
    and ignored: single-thread semantics (the reference's threads>1 branch rotates by whole
    degrees, projection.rs:94)."""
    ctx = ctx or api.default_context()
    return ctx.angle_with_projections(src_img, max_angle, step, resize_scale, threads)
