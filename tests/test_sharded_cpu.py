"""World-size-2 gloo tests of the multi-GPU host logic (no GPU): the angle-range split of the
single-huge-image mode, the bit-exact all-gather of per-angle std-devs, the exact argmax on
every rank, and the card-range split of batch mode.  The per-rank sweep is the CPU oracle
here (checker standing in for the kernel); the GPU path uses ctx.projection_sweep."""
import os
import socket

import numpy as np
import pytest
import torch.distributed as dist
import torch.multiprocessing as mp

from omr_img_corrector_b200 import api, sharded, synth


def test_shard_range_partitions():
    for n in (0, 1, 7, 300, 9000, 4096):
        for world in (1, 2, 3, 4, 8):
            cover = []
            for r in range(world):
                f, c = sharded.shard_range(n, r, world)
                assert 0 <= f <= n and 0 <= c and f + c <= n
                cover += list(range(f, f + c))
            assert cover == list(range(n)), (n, world)
    assert sharded.shard_range(9000, 3, 8) == (3375, 1125)      # BASELINE config 5
    assert sharded.shard_cards(4096, 7, 8) == (3584, 512)        # BASELINE config 4
    with pytest.raises(api.OicsError):
        sharded.shard_range(10, 2, 2)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import oracle as orc
        g, a = synth.base_card(5)
        crop = np.ascontiguousarray(g[250:650, 250:700])
        bgr = orc.rotate_mat(synth.to_bgr(crop), -2.5, 1.0, 1, orc.WHITE, 0)
        binary, s = orc.preprocess_omr(bgr, 0, 0)
        h, w = binary.shape
        N, angles, M = api.sweep_matrices(w, h, 5, 0.5, s)
        calls = []

        def sweep_fn(Ms):
            calls.append(len(Ms))
            return orc.projection_sweep(binary, Ms)

        v, hh = sharded.sweep_sharded(sweep_fn, M, rank, world)
        res = api.rule_a(hh, v, N, 0.5)
        rb = api.rule_b(v, hh, N, 0.5)
        # odd number of candidates: ragged tail shard
        v7, h7 = sharded.sweep_sharded(sweep_fn, M[:7], rank, world)
        np.savez(os.path.join(out_dir, f"r{rank}.npz"), v=v, h=hh, angle=res.angle, status=res.status, rb=rb,
                 calls=np.array(calls), v7=v7, h7=h7)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_sweep_sharded_gloo(tmp_path, orc, world):
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    g, a = synth.base_card(5)
    crop = np.ascontiguousarray(g[250:650, 250:700])
    bgr = orc.rotate_mat(synth.to_bgr(crop), -2.5, 1.0, 1, orc.WHITE, 0)
    binary, s = orc.preprocess_omr(bgr, 0, 0)
    N, angles, M = api.sweep_matrices(binary.shape[1], binary.shape[0], 5, 0.5, s)
    v, h = orc.projection_sweep(binary, M)
    ref = api.rule_a(h, v, N, 0.5)
    for r in range(world):
        R = np.load(tmp_path / f"r{r}.npz")
        assert np.array_equal(R["v"], v) and np.array_equal(R["h"], h)           # bit-exact gather
        assert (float(R["angle"]), int(R["status"])) == (ref.angle, ref.status)   # same decision on every rank
        assert float(R["rb"]) == api.rule_b(v, h, N, 0.5)
        f, c = sharded.shard_range(len(M), r, world)
        assert int(R["calls"][0]) == c                                            # each rank swept only its slice
        assert np.array_equal(R["v7"], v[:7]) and np.array_equal(R["h7"], h[:7])
    assert abs(ref.angle - 2.5) <= 0.5
