"""packages/lib/src/omr.rs."""
from . import api
from .api import OmrResult  # noqa: F401


class ResultStatus:                        # omr.rs:41-45
    Believed, NeedCheck, NotAResult = api.BELIEVED, api.NEED_CHECK, api.NOT_A_RESULT


def get_result_from_projection(src_mat, projection_max_angle, projection_angle_step, projection_max_width,
                               projection_max_height, ctx=None):
    """omr::get_result_from_projection (omr.rs:52-58)."""
    ctx = ctx or api.default_context()
    return ctx.result_from_projection(src_mat, projection_max_angle, projection_angle_step, projection_max_width,
                                      projection_max_height)


def get_result_from_edges_detection(src_mat, edges_min_line_length, edges_max_line_gap, ctx=None):
    """omr::get_result_from_edges_detection (omr.rs:231-235)."""
    ctx = ctx or api.default_context()
    return ctx.result_from_edges_detection(src_mat, edges_min_line_length, edges_max_line_gap)


def get_result_from_fourier_transform(src_mat, canny_threshold_weak, canny_threshold_strong,
                                      fourier_min_line_length, fourier_max_line_gap, ctx=None):
    """omr::get_result_from_fourier_transform (omr.rs:304-310)."""
    ctx = ctx or api.default_context()
    return ctx.result_from_fourier_transform(src_mat, canny_threshold_weak, canny_threshold_strong,
                                             fourier_min_line_length, fourier_max_line_gap)


def correct_default(src_mat, projection_max_angle, projection_angle_step, projection_max_width,
                    projection_max_height, hough_min_line_length, hough_max_line_gap, ctx=None):
    """omr::correct_default (omr.rs:339-348) on memory buffers: the caller keeps
    imread / imwrite.  -> (rotate_angle, need_check, deskewed BGR)."""
    ctx = ctx or api.default_context()
    return ctx.correct_default(src_mat, projection_max_angle, projection_angle_step, projection_max_width,
                               projection_max_height, hough_min_line_length, hough_max_line_gap)
