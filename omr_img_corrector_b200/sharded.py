"""Multi-GPU drivers of the projection method.

* ``shard_cards``        -- card batches (BASELINE configs 1-4): contiguous card ranges per
                            rank, no collective on the data path (the reference's analogue
                            is one OS thread per card, thread_pool.rs:56-73).
* ``sweep_sharded``      -- single-huge-image mode (config 5): every rank holds the image and
                            evaluates a contiguous slice of the candidate angles; the
                            per-angle (v_sd, h_sd) slices are all-gathered (NCCL on GPUs,
                            gloo in the CPU tests) and EVERY rank then applies the exact
                            scalar argmax rule to the full arrays.  The rule is not a plain
                            max (strict-< with tie bookkeeping, omr.rs:147-221), so the
                            exchange is gather + local reduce, not an arg-max all-reduce.

torch.distributed is plumbing only: the payload is 16 bytes per angle.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib, api


def shard_range(n: int, rank: int, world: int):
    """-> (first, count): oics_shard_range."""
    f, c = C.c_int32(), C.c_int32()
    _lib.check(_lib.load().oics_shard_range(int(n), int(rank), int(world), C.byref(f), C.byref(c)))
    return f.value, c.value


def shard_cards(n_cards: int, rank: int, world: int):
    """Contiguous card range of `rank` (same split rule as the angle shards)."""
    return shard_range(n_cards, rank, world)


def all_gather_blocks(sd, rank: int, world: int, group=None):
    """In-place all-gather of the exchange buffer sd[world][2][per] (oics_shard_layout): every rank has
    filled its own block sd[rank]; afterwards every rank holds all blocks.  `sd` is a torch tensor --
    on the GPU the collective is ONE ncclAllGather of 2 * per doubles per rank over NVLink, on the CPU
    (the gloo tests) the same call.  float64 travels bit-exactly (no reduction arithmetic)."""
    import torch.distributed as dist
    if world > 1:
        flat = sd.reshape(-1)
        blk = flat.numel() // world
        dist.all_gather_into_tensor(flat, flat[rank * blk:(rank + 1) * blk].clone(), group=group)
    return sd


def all_gather_slices(v_local, h_local, n_total: int, rank: int, world: int, group=None, device=None):
    """All-gathers per-rank (v_sd, h_sd) slices into full [n_total] arrays on every rank, through the
    library's exchange layout (block = slice of v_sd, slice of h_sd, each padded to ceil(n / world))."""
    import torch

    per = api.shard_layout(n_total, world)
    first, count = shard_range(n_total, rank, world)
    assert len(v_local) == count and len(h_local) == count
    sd = torch.zeros((world, 2, per), dtype=torch.float64)
    sd[rank, 0, :count] = torch.from_numpy(np.ascontiguousarray(v_local, np.float64))
    sd[rank, 1, :count] = torch.from_numpy(np.ascontiguousarray(h_local, np.float64))
    if device is not None:
        sd = sd.to(device)
    all_gather_blocks(sd, rank, world, group)
    return api.shard_unpack(sd.cpu().numpy(), n_total, world)


def sweep_sharded(sweep_fn, M, rank: int, world: int, group=None, device=None):
    """Evaluates this rank's slice of the forward matrices M [n, 6] with
    ``sweep_fn(M_slice) -> (v_sd, h_sd)`` and returns the gathered full arrays."""
    M = np.ascontiguousarray(M, np.float64).reshape(-1, 6)
    first, count = shard_range(len(M), rank, world)
    if count:
        v, h = sweep_fn(M[first:first + count])
    else:
        v = h = np.zeros(0, np.float64)
    return all_gather_slices(v, h, len(M), rank, world, group, device)


def sweep_sharded_device(ctx: api.Context, binary, M, rank: int, world: int, group=None, device=None, sd=None):
    """Config 5 on GPUs: oics_sweep_sharded writes this rank's std-dev slices straight into its block of a
    DEVICE exchange buffer, NCCL all-gathers the buffer in place, and the unpacked arrays come back to
    the host (144 KB for 9000 angles).  -> (v_sd[n], h_sd[n]), identical on every rank."""
    import torch
    M = np.ascontiguousarray(M, np.float64).reshape(-1, 6)
    per = api.shard_layout(len(M), world)
    if sd is None:
        sd = torch.zeros((world, 2, per), dtype=torch.float64, device=device)
    ctx.sweep_sharded(binary, M, rank, world, sd)
    all_gather_blocks(sd, rank, world, group)
    return api.shard_unpack(sd.cpu().numpy(), len(M), world)


def result_from_projection_sharded(ctx: api.Context, binary, max_angle: int, step: float, rot_scale: float,
                                   rank: int, world: int, group=None, device=None, rule: str = "a"):
    """Config 5: the projection estimate of one (huge) pre-processed binary image with the
    angle range split over `world` GPUs.  `binary` may be a numpy array or a device image.
    -> OmrResult (rule 'a', omr.rs:147-221) or angle (rule 'b', projection.rs:125-190);
    identical on every rank."""
    p, w, h, ch, stride, keep = api._describe(binary)
    N, angles, M = api.sweep_matrices(w, h, max_angle, step, rot_scale)
    if device is not None:
        v, hh = sweep_sharded_device(ctx, binary, M, rank, world, group, device)
    else:
        v, hh = sweep_sharded(lambda Ms: ctx.projection_sweep(binary, Ms), M, rank, world, group, device)
    if rule == "a":
        return api.rule_a(hh, v, N, step)
    return api.rule_b(v, hh, N, step)
