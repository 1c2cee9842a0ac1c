"""CPU tests: the C-ABI library loads, exports every symbol the header declares, and its
host-side scalar logic (geometry, argmax rules, segment-angle rules, reconcile) agrees
with the oracle.  No compute entry is called (no GPU here)."""
import math
import os
import re

import numpy as np
import pytest

from omr_img_corrector_b200 import _lib, api

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "oics_cuda.h")).read()
    declared = set(re.findall(r"OICS_API\s+[\w\s\*]+?\b(oics_\w+)\s*\(", hdr))
    assert len(declared) >= 35
    lib = _lib.load()
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in the header but not exported"
    assert declared == set(_lib.SIGNATURES), declared ^ set(_lib.SIGNATURES)


def test_no_gpu_means_loud_failure():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(api.OicsError) as e:
        api.Context(0)
    assert e.value.code == _lib.ERR_CUDA and "no CPU fallback" in str(e.value)


def test_probes_reject_a_null_context_without_touching_cuda():
    import ctypes as C
    lib, g = _lib.load(), C.c_double()
    assert lib.oics_probe_l2_read(None, 32 << 20, 4, C.byref(g)) == _lib.ERR_INVALID_ARG
    assert lib.oics_probe_smem_read(None, 16, C.byref(g)) == _lib.ERR_INVALID_ARG
    assert b"invalid argument" in lib.oics_last_error()


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "omr_img_corrector_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cpp", ".h")):
                src = open(os.path.join(dirpath, f), errors="replace").read()
                assert "oracle" not in src.lower(), f"{f} mentions the oracle"


def test_angle_grid_is_half_open(orc):
    for max_angle, step, expect in [(15, 0.1, 150), (45, 0.2, 225), (45, 0.01, 4500), (0, 0.1, 0), (45, 1e-9, 65535)]:
        assert api.angle_half_range(max_angle, step) == expect
        assert orc.sweep_angles(max_angle, step)[0] == expect
    N, angles, M = api.sweep_matrices(2480, 3508, 15, 0.1, 1.0)
    assert N == 150 and len(angles) == 300 and angles[0] == -15.0 and angles[153] == 3 * 0.1
    _, _, Mo = orc.sweep_matrices(2480, 3508, 15, 0.1, 1.0)
    assert np.array_equal(M, Mo)
    _, _, M2 = api.sweep_matrices(248, 230, 45, 0.2, 0.2)
    assert np.array_equal(M2, orc.sweep_matrices(248, 230, 45, 0.2, 0.2)[2])


def test_rotate_geometry(orc):
    for (w, h) in [(2480, 3508), (1240, 1150), (1237, 1300), (7, 5)]:
        for ang in [-15.0, -2.8, 0.0, 0.1, 3.7, 44.9, 90.0, -135.5]:
            for clip in (0, 1):
                M, dw, dh = api.rotate_geometry(w, h, ang, 1.0, clip)
                Mo, dwo, dho = orc.rotate_geometry(w, h, ang, 1.0, clip)
                assert (dw, dh) == (dwo, dho) and np.array_equal(M, Mo)


def test_projection_scale(orc):
    for args in [(1240, 1150, 248, 230), (1237, 1300, 248, 230), (2480, 3508, 0, 0), (100, 100, 300, -1)]:
        assert api.projection_scale(*args) == orc.projection_scale(*args)


def _rand_sd(rng, n, ties):
    v = rng.random(n) * 10
    h = rng.random(n) * 10
    if ties:
        k = rng.integers(0, n, size=max(2, n // 4))
        h[k] = h.max() if ties > 1 else h[k[0]]
        if ties > 2:
            v[k] = v.max()
    return v, h


def test_rule_a_matches_oracle(orc):
    rng = np.random.default_rng(0)
    for trial in range(300):
        N = int(rng.integers(1, 40))
        v, h = _rand_sd(rng, 2 * N, trial % 4)
        if trial % 7 == 0:
            h[:] = 0.0
            v[:] = 0.0
        got = api.rule_a(h, v, N, 0.1)
        ref = orc.rule_a(h, v, N, 0.1)
        assert got.status == ref.status and got.angle == ref.angle and got.candidates == ref.candidates


def test_rule_a_literal_python(orc):
    """omr.rs:147-221 transcribed in Python as a third opinion."""
    rng = np.random.default_rng(1)
    for trial in range(200):
        N = int(rng.integers(1, 30))
        v, h = _rand_sd(rng, 2 * N, trial % 4)
        mh = mv = 0.0
        ph = pv = 1
        most = []
        for i, deg in enumerate(range(-N, N)):
            a = float(deg) * 0.2
            if mh < h[i]:
                mh, mv, ph, pv, most = h[i], v[i], 1, 1, [a]
            elif mh == h[i]:
                ph += 1
                if mv < v[i]:
                    pv, mv, most = 1, v[i], [a]
                elif mv == v[i]:
                    pv += 1
                    most.append(a)
        if ph == 1 and pv == 1:
            exp = (most[0], 0)
        elif len(most) == 1:
            exp = (most[0], 1)
        else:
            exp = (0.0, 2)
        got = api.rule_a(h, v, N, 0.2)
        assert (got.angle, got.status) == exp and got.candidates == most


def test_rule_b_matches_oracle(orc):
    rng = np.random.default_rng(2)
    for trial in range(300):
        N = int(rng.integers(1, 40))
        v, h = _rand_sd(rng, 2 * N, trial % 4)
        if trial % 5 == 0:
            i = int(rng.integers(0, 2 * N))
            v[i] = 100.0
            h[i] = 100.0
        if trial % 11 == 0:
            v[:] = 0
            h[:] = 0
        assert api.rule_b(v, h, N, 0.2) == orc.rule_b(v, h, N, 0.2)


def _rand_lines(rng, n):
    base = rng.uniform(-40, 40)
    lines = []
    for _ in range(n):
        a = math.radians(base + rng.normal(0, 0.05) if rng.random() < 0.6 else rng.uniform(-180, 180))
        L = rng.uniform(50, 600)
        x1, y1 = rng.uniform(0, 1000, 2)
        lines.append([int(x1), int(y1), int(x1 + L * math.cos(a)), int(y1 + L * math.sin(a))])
    return np.array(lines, np.int32).reshape(-1, 4)


def test_lines_angle_variants_match_oracle(orc):
    rng = np.random.default_rng(3)
    for trial in range(100):
        lines = _rand_lines(rng, int(rng.integers(1, 60)))
        got = api.lines_angle(lines, api.LINES_OMR)
        ref = orc.lines_angle_omr(lines)
        assert got.angle == ref.angle and got.status == ref.status and got.candidates == ref.candidates
        assert api.lines_angle(lines, api.LINES_HOUGH).angle == orc.lines_angle_hough(lines)
        assert api.lines_angle(lines, api.LINES_FFT).angle == orc.lines_angle_fft(lines)


def test_lines_angle_empty_behaviour(orc):
    empty = np.zeros((0, 4), np.int32)
    for variant in (api.LINES_OMR, api.LINES_HOUGH):
        with pytest.raises(api.OicsError) as e:
            api.lines_angle(empty, variant)
        assert e.value.code == _lib.ERR_NO_LINES          # the reference panics (omr.rs:272, hough.rs:74)
    assert api.lines_angle(empty, api.LINES_FFT).angle == 0.0 == orc.lines_angle_fft(empty)


def test_reconcile_matches_oracle(orc):
    rng = np.random.default_rng(4)
    for trial in range(300):
        status = trial % 3
        cand = [float(x) for x in np.round(rng.uniform(-5, 5, int(rng.integers(0, 5))), 1)]
        if status != 2 and not cand:
            cand = [1.0]
        p = api.OmrResult(cand[0] if status != 2 else 0.0, status, cand)
        e = float(np.round(rng.uniform(-5, 5), 2)) if trial % 4 else (cand[0] + 0.03 if cand else 0.0)
        assert api.reconcile(p, e) == orc.reconcile(orc.OmrResult(p.angle, status, cand), e)


def test_hough_dims(orc):
    for (w, h) in [(2480, 3508), (1240, 1150), (310, 288)]:
        assert api.hough_dims(w, h) == orc.hough_dims(w, h)
    assert api.hough_dims(2480, 3508) == (180, 11977)


def test_rule_a_batch_matches_rule_a():
    from omr_img_corrector_b200 import api
    rng = np.random.default_rng(5)
    N = 20
    h = rng.integers(0, 6, (7, 2 * N)).astype(np.float64)       # small alphabet -> exact ties
    v = rng.integers(0, 6, (7, 2 * N)).astype(np.float64)
    ang, stat, ncand = api.rule_a_batch(h, v, N, 0.5)
    for c in range(7):
        r = api.rule_a(h[c], v[c], N, 0.5)
        assert (ang[c], stat[c], ncand[c]) == (r.angle, r.status, len(r.candidates))


def test_task_events_follow_the_reference_vocabulary():
    """task.rs:19-68: start_running_task, then task_completed with finished | debatable | error."""
    from omr_img_corrector_b200 import task

    class FakeCtx:
        def correct_batch(self, cards, *params, outs=None):
            n = len(cards)
            ang = np.arange(n, dtype=np.float64)
            need = np.array([i % 3 == 1 for i in range(n)])
            stat = np.array([-3 if i % 3 == 2 else 0 for i in range(n)], np.int32)
            sizes = np.tile(np.array([[4, 2]], np.int32), (n, 1))
            if outs is not None:
                for i, o in enumerate(outs):
                    o[:24] = i
            return ang, need, stat, sizes

    tasks = [task.Task(10 + i, np.zeros((5, 6, 3), np.uint8), f"out{i}.jpg") for i in range(4)]
    tasks.append(task.Task(99, np.zeros((5, 6), np.uint8), "gray.jpg"))           # not a BGR card -> error
    events, written = [], {}
    task.run_tasks(FakeCtx(), tasks, 45, 0.2, 248, 230, 150.0, 50.0, lambda e, p: events.append((e, p)),
                   write=lambda path, img: written.__setitem__(path, img.copy()))
    done = {p["task_id"]: p for e, p in events if e == "task_completed"}
    started = [p["task_id"] for e, p in events if e == "start_running_task"]
    assert sorted(started) == [10, 11, 12, 13, 99]
    assert [done[k]["result"] for k in (10, 11, 12, 13, 99)] == ["finished", "debatable", "error", "finished", "error"]
    assert done[11]["output_path"] == "out1.jpg"
    assert set(written) == {"out0.jpg", "out1.jpg", "out3.jpg"} and written["out3.jpg"].shape == (2, 4, 3)
    for tid in (10, 11, 12, 13, 99):                       # a task starts before it completes
        i0 = next(i for i, (e, p) in enumerate(events) if e == "start_running_task" and p["task_id"] == tid)
        i1 = next(i for i, (e, p) in enumerate(events) if e == "task_completed" and p["task_id"] == tid)
        assert i0 < i1


def test_shard_layout_and_unpack():
    """Config 5's gather layout sd[world][2][per] (oics_shard_layout / oics_shard_unpack)."""
    from omr_img_corrector_b200 import sharded
    for n, world in [(9000, 8), (9000, 7), (5, 8), (1, 1), (0, 3), (300, 2)]:
        per = api.shard_layout(n, world)
        assert per == (n + world - 1) // world
        sd = np.zeros((world, 2, max(per, 1)))[:, :, :per]
        sd = np.ascontiguousarray(sd)
        v0, h0 = np.arange(n, dtype=np.float64), -np.arange(n, dtype=np.float64) - 0.5
        for r in range(world):
            f, c = sharded.shard_range(n, r, world)
            sd[r, 0, :c] = v0[f:f + c]
            sd[r, 1, :c] = h0[f:f + c]
        v, h = api.shard_unpack(sd, n, world)
        assert np.array_equal(v, v0) and np.array_equal(h, h0)


def test_rule_a_empty_range_mirrors_the_panic():
    """omr.rs:210-211: with N == 0 the reference indexes an empty Vec and panics."""
    import pytest
    with pytest.raises(api.OicsError) as e:
        api.rule_a(np.zeros(0), np.zeros(0), 0, 0.2)
    assert e.value.code == -1


def test_transformable_matrix_host_members():
    """transfer::TransformableMatrix (transfer.rs:16-277): construction, deep copies, accessors (no device)."""
    from omr_img_corrector_b200 import transfer
    img = np.arange(24, dtype=np.uint8).reshape(2, 4, 3)
    tm = transfer.TransformableMatrix.from_matrix(img)
    img[0, 0, 0] = 99
    assert tm.get_mat()[0, 0, 0] == 0                                  # from_matrix deep-copies (transfer.rs:55-59)
    c = tm.clone()
    c.matrix[0, 0, 0] = 7
    assert tm.get_mat()[0, 0, 0] == 0 and tm.get_bytes() == bytes(range(24))
    assert tm.scale_self(1.0) is tm                                    # the early-out of transfer.rs:67-69
    assert transfer.TransformableMatrix.default().get_mat().size == 0


def test_bench_reference_arm_sample(orc):
    """bench.py's CPU arm: card-parallel workers, file to file, sweep bounded to a sample of the 300 angles; the
    two rates it reports are the sweep-only region (`value`) and the whole of correct_default with codecs (`e2e`)."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("bench", os.path.join(ROOT, "bench.py"))
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    ref = bench.CpuReference(threads=2, n_sample=2)
    r = ref.step()
    assert r["sweep_cards_per_s"] > r["e2e_cards_per_s"] > 0
    assert abs(r["sweep_cards_per_s"] - 2 / (300 * r["t_angle_s"])) < 1e-9
    per_card = r["t_decode_s"] + r["t_pre_s"] + 300 * r["t_angle_s"] + r["t_warp_s"] + r["t_encode_s"]
    assert abs(r["e2e_cards_per_s"] - 2 / per_card) < 1e-9
    assert "300 sweep angles" in ref.sample_text()
    cfg_a = bench.workload_config(512, 1, 16, 128)
    assert cfg_a == bench.workload_config(512, 1, 16, 128) and "workload" in cfg_a     # both arms print the same config


def test_work_queue_default_respects_the_host(tmp_path):
    """Importing the package asks for one hardware work queue per stream (the file path keeps ~24 busy) unless the
    host exported its own CUDA_DEVICE_MAX_CONNECTIONS (process environment)."""
    import subprocess
    import sys
    code = "import os, omr_img_corrector_b200; print(os.environ['CUDA_DEVICE_MAX_CONNECTIONS'])"
    env = {k: v for k, v in os.environ.items() if k != "CUDA_DEVICE_MAX_CONNECTIONS"}
    env["PYTHONPATH"] = ROOT
    assert subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, check=True).stdout.strip() == "32"
    env["CUDA_DEVICE_MAX_CONNECTIONS"] = "4"
    assert subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, check=True).stdout.strip() == "4"
