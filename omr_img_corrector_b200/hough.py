"""packages/lib/src/hough.rs."""
from . import api


def get_angle_with_hough(gray_tm, min_line_length, max_line_gap, file_name=None, edge_image_output_dir=None,
                         ctx=None):
    """hough::get_angle_with_hough (hough.rs:17-23).  The debug edge image the reference
    writes (hough.rs:45-63, 91-96) is not produced; file_name / edge_image_output_dir are
    accepted for signature compatibility.  Raises OicsError(NO_LINES) where the reference
    panics on an empty segment list (hough.rs:74)."""
    ctx = ctx or api.default_context()
    return ctx.angle_with_hough(gray_tm, min_line_length, max_line_gap)
