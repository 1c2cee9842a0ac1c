"""oics_projection_standard_deviations_static: the allocation-free variant of
oics_projection_standard_deviations (scratch in a file-scope static __device__ array, noalloc.cu), called
through the C ABI and compared with the existing entry point -- bit for bit, the tolerance the other std-dev
parity tests would allow being SD_RTOL = 1e-12 -- on calls that follow each other and share the array.
Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

from omr_img_corrector_b200 import _lib, api

pytestmark = pytest.mark.gpu

MAX_SIDE = 16384


def _card(seed, h, w):
    """A binary card (0 / 255) with row and column structure, so the two std-devs differ."""
    rng = np.random.default_rng(seed)
    img = np.where(rng.random((h, w)) < 0.03, 0, 255).astype(np.uint8)
    for y in range(20, h - 20, 37):
        img[y:y + 5, w // 8: w - w // 8] = 0
    for x in range(15, w - 15, 53):
        img[h // 6: h - h // 6, x:x + 3] = 0
    return img


def _static(torch, d_img, d_sd):
    api.projection_standard_deviations_static(d_img, d_sd, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    return tuple(d_sd.cpu().tolist())


def test_static_scratch_entry_equals_the_allocating_one():
    import torch
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    d_sd = torch.zeros(2, dtype=torch.float64, device=dev)
    # the big card first, the small ones after it: whatever the first call left in the array must not leak
    cases = [(0, 3508, 2480, 0), (1, 301, 417, 0), (2, 640, 1027, 5), (3, 3508, 2480, 0), (4, 1, 1, 0), (5, 7, 4100, 0)]
    got, imgs = [], []
    for k, (seed, h, w, pad) in enumerate(cases):
        img = _card(seed, h, w) if h > 40 and w > 40 else np.zeros((h, w), np.uint8) + (255 if seed == 4 else 0)
        if h == 7:
            img[::2, ::3] = 255
        d_full = torch.from_numpy(np.pad(img, ((0, 0), (0, pad)), constant_values=0)).to(dev)
        d_img = d_full[:, :w]                      # stride w + pad: the padding must not be counted
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info(0)[0]
        got.append(_static(torch, d_img, d_sd))
        if k >= 1:                                  # (the first call may load the kernels' module)
            assert torch.cuda.mem_get_info(0)[0] == free_before, "the static entry point changed the free device memory"
        imgs.append(img)
        del d_full, d_img
    # two calls in a row on the same image: identical
    d_img = torch.from_numpy(imgs[0]).to(dev)
    a = _static(torch, d_img, d_sd)
    b = _static(torch, d_img, d_sd)
    assert a == b == got[0]
    # against the existing entry point
    ctx = api.Context(0)
    try:
        for img, g in zip(imgs, got):
            ref = ctx.projection_standard_deviations(img)
            assert g == tuple(ref), (img.shape, g, ref)
    finally:
        ctx.close()


def test_static_scratch_entry_rejects_what_it_cannot_hold():
    import torch
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    lib = _lib.load()
    d_sd = torch.zeros(2, dtype=torch.float64, device=dev)
    d_img = torch.zeros((4, MAX_SIDE + 1), dtype=torch.uint8, device=dev)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = lib.oics_projection_standard_deviations_static(C.c_void_p(d_img.data_ptr()), MAX_SIDE + 1, 4, MAX_SIDE + 1,
                                                        C.c_void_p(d_sd.data_ptr()), st)
    assert rc == -7 and b"16384" in lib.oics_last_error()
    rc = lib.oics_projection_standard_deviations_static(C.c_void_p(d_img.data_ptr()), 4, MAX_SIDE + 1, 4,
                                                        C.c_void_p(d_sd.data_ptr()), st)
    assert rc == -7
    # host memory is refused: nothing is staged
    host = np.zeros((8, 8), np.uint8)
    rc = lib.oics_projection_standard_deviations_static(C.c_void_p(host.ctypes.data), 8, 8, 8, C.c_void_p(d_sd.data_ptr()), st)
    assert rc == -1
    # the widest image it does hold
    wide = torch.full((3, MAX_SIDE), 255, dtype=torch.uint8, device=dev)
    wide[1, 5:9] = 0
    api.projection_standard_deviations_static(wide, d_sd, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    ctx = api.Context(0)
    try:
        assert tuple(d_sd.cpu().tolist()) == tuple(ctx.projection_standard_deviations(wide.cpu().numpy()))
    finally:
        ctx.close()
