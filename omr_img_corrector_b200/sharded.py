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


def all_gather_slices(v_local, h_local, n_total: int, rank: int, world: int, group=None, device=None):
    """All-gathers per-rank (v_sd, h_sd) slices into full [n_total] arrays on every rank.

    Slices are padded to ceil(n/world) so one fixed-size all_gather suffices; float64
    payload travels bit-exactly (no reduction arithmetic is involved)."""
    import torch
    import torch.distributed as dist

    per = (n_total + world - 1) // world
    first, count = shard_range(n_total, rank, world)
    assert len(v_local) == count and len(h_local) == count
    buf = torch.zeros((2, per), dtype=torch.float64)
    buf[0, :count] = torch.from_numpy(np.ascontiguousarray(v_local, np.float64))
    buf[1, :count] = torch.from_numpy(np.ascontiguousarray(h_local, np.float64))
    if device is not None:
        buf = buf.to(device)
    if world > 1:
        out = torch.empty(world * 2 * per, dtype=torch.float64, device=buf.device)
        dist.all_gather_into_tensor(out, buf.reshape(-1), group=group)
        out = out.reshape(world, 2, per)
    else:
        out = buf[None]
    out = out.cpu().numpy()
    v = np.concatenate([out[r, 0] for r in range(world)])[:n_total]
    h = np.concatenate([out[r, 1] for r in range(world)])[:n_total]
    return np.ascontiguousarray(v), np.ascontiguousarray(h)


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


def result_from_projection_sharded(ctx: api.Context, binary, max_angle: int, step: float, rot_scale: float,
                                   rank: int, world: int, group=None, device=None, rule: str = "a"):
    """Config 5: the projection estimate of one (huge) pre-processed binary image with the
    angle range split over `world` GPUs.  `binary` may be a numpy array or a device image.
    -> OmrResult (rule 'a', omr.rs:147-221) or angle (rule 'b', projection.rs:125-190);
    identical on every rank."""
    p, w, h, ch, stride, keep = api._describe(binary)
    N, angles, M = api.sweep_matrices(w, h, max_angle, step, rot_scale)
    v, hh = sweep_sharded(lambda Ms: ctx.projection_sweep(binary, Ms), M, rank, world, group, device)
    if rule == "a":
        return api.rule_a(hh, v, N, step)
    return api.rule_b(v, hh, N, step)
