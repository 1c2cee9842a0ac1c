"""packages/lib/src/fft.rs."""
from . import api


def get_fft_image(gray_tm, ctx=None):
    """fft::get_fft_image (fft.rs:124) -> (magnitude_image, magnitude_log_image)."""
    ctx = ctx or api.default_context()
    return ctx.fft_image(gray_tm, want_magnitude=True)


def get_angle_with_fft(gray_tm, canny_threshold_1, canny_threshold_2, min_line_length, max_line_gap,
                       file_name=None, edge_image_output_dir=None, ctx=None):
    """fft::get_angle_with_fft (fft.rs:145-154), without the debug image."""
    ctx = ctx or api.default_context()
    return ctx.angle_with_fft(gray_tm, canny_threshold_1, canny_threshold_2, min_line_length, max_line_gap)
