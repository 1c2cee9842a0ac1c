"""B200-native deskew hot path of ch1ny/omr-img-corrector (crate `oics`).

The package holds only what the path needs: `csrc/` (hand-written sm_100a CUDA
kernels + the C ABI of include/oics_cuda.h, built into liboics_cuda.so) and a
host-side mirror of the reference's module layout -- `omr`, `projection`,
`hough`, `fft`, `transfer`, `calculate`, `types` (and `task`, the Tauri batch path) -- whose functions keep the
reference's names, argument meaning and error behaviour and call the C ABI.
Importing the package does not need a GPU; calling a compute function does.
"""
from . import _lib                                   # noqa: F401
from .api import (BELIEVED, CLIP_CONTAIN, CLIP_DEFAULT, INTER_LINEAR, INTER_NEAREST, NEED_CHECK,  # noqa: F401
                  NOT_A_RESULT, WARP_POLAR_LINEAR, WHITE, Context, DeviceImage, OmrResult, OicsError,
                  default_context)

__all__ = ["Context", "DeviceImage", "OmrResult", "OicsError", "default_context", "omr", "projection", "hough",
           "fft", "transfer", "calculate", "types", "task"]
