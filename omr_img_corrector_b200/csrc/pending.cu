// Entry points whose kernels are not written yet report OICS_ERR_UNIMPLEMENTED
// (never a CPU fallback).
#include "common.h"

namespace oics {

int32_t fft_image_device(oics_ctx *, const uint8_t *, int, int, size_t, uint8_t *, uint8_t *)
{
    set_error("fft image kernel not implemented yet");
    return OICS_ERR_UNIMPLEMENTED;
}
int32_t canny_device(oics_ctx *, const uint8_t *, int, int, int, size_t, int, int, uint8_t *)
{
    set_error("canny kernel not implemented yet");
    return OICS_ERR_UNIMPLEMENTED;
}
int32_t hough_accumulate_device(oics_ctx *, const uint8_t *, int, int, size_t, double, double, int32_t *)
{
    set_error("hough accumulate kernel not implemented yet");
    return OICS_ERR_UNIMPLEMENTED;
}
int32_t hough_lines_p_device(oics_ctx *, const uint8_t *, int, int, size_t, double, double, int, int, int, int32_t *, int,
                             int *)
{
    set_error("hough lines kernel not implemented yet");
    return OICS_ERR_UNIMPLEMENTED;
}

}  // namespace oics
