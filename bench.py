#!/usr/bin/env python
"""Headline benchmark: cards/sec deskewed (A4@300dpi, +-15 deg at 0.1 deg) on N B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A "step" is one pass of the hot path over one batch of synthetic cards:
  value : BASELINE config 4 shard -- CARDS pre-processed 8-bit binary A4 cards per GPU,
          resident in HBM; fused projection sweep (300 candidate angles) + rule-A argmax
          per card (omr.rs:147-221).  Whole-job cards/s over all ranks.  P0 (pre-processing)
          and W1 (the final warp) are NOT in this region; e2e has them.
  e2e   : omr::correct_default as the reference runs it, FILE TO FILE, through the C ABI
          (oics_correct_batch_files): JPEG files in pinned host memory in (imread, omr.rs:349),
          decode + P0 + sweep + rule A (+ Hough fallback) + CONTAIN warp + JPEG quality 100 encode
          (imwrite, omr.rs:447-448) on the device, files copied back.  Host<->device copies are
          inside the timed region.  Side fields: the same with raw BGR pixels in and out
          (e2e.raw_pixels) and raw pixels in / JPEG out (e2e.jpeg_out).
  configs : driver-visible numbers for BASELINE configs 2 (Hough), 3 (FFT), 5 (one 9921x14032 scan,
          +-45 deg at 0.01 deg, the angle range sharded over the N ranks through oics_sweep_sharded +
          one NCCL all-gather) and the reference's shipped operating point (45, 0.2, 248, 230, 150, 50),
          each checked against the oracle OUTSIDE its timed region.
  --impl reference : the reference's CPU path (the oracle restatement of its OpenCV call
          sequence; the Rust crate itself cannot be built here) on the host cores, card-parallel,
          codecs included in its e2e.
Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")   # see omr_img_corrector_b200/_lib.py: one work queue per decode stream
sys.path.insert(0, ROOT)

W, H = 2480, 3508
MAX_ANGLE, STEP = 15, 0.1
N_ANGLES = 300
METRIC = "cards/sec deskewed (A4@300dpi, +-15deg/0.1deg)"
UNIT = "cards/s"
DEFAULT_OP = (45, 0.2, 248, 230, 150.0, 50.0)     # lib.rs:92-97, getLibParams.ts:29-34


def workload_config(cards_per_gpu, n_gpus, distinct, e2e_cards):
    return {
        "workload": "BASELINE config 4 shard: synthetic A4@300dpi (2480x3508) OMR cards, projection method, "
                    "+-15deg at 0.1deg (300 candidate angles), rule-A argmax",
        "cards_per_gpu_per_step": cards_per_gpu,
        "global_cards_per_step": cards_per_gpu * n_gpus,
        "distinct_cards_per_gpu": distinct,
        "angles": N_ANGLES,
        "sharding": f"cards sharded over {n_gpus} GPU(s), no collective on the data path",
        "l2": "inputs larger than L2 (cards_per_gpu x 8.7 MB binary cards streamed from HBM each step)",
        "value_region": "sweep + rule-A argmax on HBM-resident pre-processed cards; pre-processing (P0) and the final "
                        "warp (W1) are outside it (they are inside e2e)",
        "e2e_workload": f"omr::correct_default file to file via oics_correct_batch_files: {e2e_cards} JPEG (quality 100) "
                        "files per GPU per step in pinned host memory in, JPEG files out; decode, P0, sweep, rule A, "
                        "CONTAIN warp and encode on the device (e2e.value).  e2e.raw_pixels: BGR pixels in and out "
                        "(oics_correct_batch); e2e.jpeg_out: BGR pixels in, JPEG files out (oics_correct_batch_jpeg)",
    }


# --------------------------------------------------------------------------- clocks
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.index)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
        sm = [float(r[0]) for r in self.rows if len(r) >= 7 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 7 and r[1].replace(".", "").isdigit()]
        reasons = set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            if len(r) >= 7:
                for k, nm in enumerate(names):
                    if r[3 + k].lower().startswith("active"):
                        reasons.add(nm)
        # the busiest half of the samples is "under load"
        sm_sorted = sorted(sm)
        load = sm_sorted[len(sm_sorted) // 2:] if sm_sorted else []
        return {"sm_mhz": float(np.median(load)) if load else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# --------------------------------------------------------------------------- CPU arm
def host_cores():
    """-> (threads this process may run on, physical cores of the box or None)."""
    try:
        threads = len(os.sched_getaffinity(0))
    except AttributeError:
        threads = os.cpu_count() or 1
    phys = None
    try:
        import psutil
        phys = psutil.cpu_count(logical=False)
    except Exception:
        pass
    return max(1, threads), phys


class CpuReference:
    """The reference's CPU path (oracle port), CARD-PARALLEL like its thread pool (thread_pool.rs:41-73): one
    worker thread per host thread, each running omr::correct_default on its own card from file bytes to file
    bytes -- imread (omr.rs:349), pre-processing, the sweep, the final CONTAIN warp, imwrite(JPEG, 100)
    (omr.rs:447-448) -- with the sweep BOUNDED to `n_sample` of its 300 angles evenly spaced (every angle is
    the same work: a full-resolution NN warpAffine into a fresh image + zero counts + two std-devs).  The
    per-stage times are the means over the workers while all of them are busy, so memory-bandwidth contention
    between cores is in the numbers.
        sweep-only cards/s (the region `value` times)  = threads / (300 t_angle)
        file-to-file cards/s (the region `e2e` times)  = threads / (t_decode + t_pre + 300 t_angle + t_warp + t_encode)
    """

    def __init__(self, threads, n_sample=8):
        from omr_img_corrector_b200 import synth
        from oracle import oracle as orc
        self.orc, self.threads, self.n_sample = orc, threads, n_sample
        self.files, self.skews = [], []
        for seed in (0, 1):
            g, a = synth.base_card(seed)
            bgr = orc.rotate_mat(synth.to_bgr(g), -a, 1.0, 1, orc.WHITE, 0)
            self.files.append(orc.jpeg_encode(bgr, 100))
            self.skews.append(a)
        N, self.angles, self.Ms = orc.sweep_matrices(W, H, MAX_ANGLE, STEP, 1.0)
        self.idx = np.unique(np.linspace(0, len(self.Ms) - 1, n_sample).round().astype(int))

    def _worker(self, k, out):
        orc = self.orc
        t = [time.perf_counter()]
        bgr = orc.jpeg_decode(self.files[k % len(self.files)])
        t.append(time.perf_counter())
        binary, s = orc.preprocess_omr(bgr, 0, 0)
        t.append(time.perf_counter())
        orc.projection_sweep(binary, self.Ms[self.idx], threads=1)
        t.append(time.perf_counter())
        rot = orc.rotate_mat(bgr, self.skews[k % len(self.files)], 1.0, 0, orc.WHITE, 1)
        t.append(time.perf_counter())
        orc.jpeg_encode(rot, 100)
        t.append(time.perf_counter())
        out[k] = np.diff(t)

    def step(self):
        out = [None] * self.threads
        ts = [threading.Thread(target=self._worker, args=(k, out)) for k in range(self.threads)]
        t0 = time.perf_counter()
        [t.start() for t in ts]
        [t.join() for t in ts]
        wall = time.perf_counter() - t0
        d = np.mean(np.stack(out), axis=0)
        t_dec, t_pre, t_sw, t_warp, t_enc = [float(x) for x in d]
        t_angle = t_sw / len(self.idx)
        per_card = t_dec + t_pre + N_ANGLES * t_angle + t_warp + t_enc
        return {
            "sweep_cards_per_s": self.threads / (N_ANGLES * t_angle),
            "e2e_cards_per_s": self.threads / per_card,
            "angle_evals_per_s": self.threads / t_angle,
            "t_decode_s": t_dec, "t_pre_s": t_pre, "t_angle_s": t_angle, "t_warp_s": t_warp, "t_encode_s": t_enc,
            "wall_s": wall, "cpu_seconds": float(np.sum(np.stack(out))),
        }

    def sample_text(self):
        return (f"{self.threads} worker threads, one A4 card each (card-parallel): imread of its JPEG, pre-processing, "
                f"{len(self.idx)} of the 300 sweep angles evenly spaced, the final CONTAIN warp, imwrite(JPEG, 100); "
                "value = threads / (300 t_angle) (the sweep + argmax region), e2e = threads / (t_decode + t_pre + "
                "300 t_angle + t_warp + t_encode), per-stage means over the concurrently busy workers")


def run_reference(args, rank, world):
    if rank != 0:
        return
    threads, phys = host_cores()
    ref = CpuReference(threads)
    for _ in range(args.warmup):
        ref.step()
    rates, t0 = [], time.perf_counter()
    for _ in range(args.steps):
        rates.append(ref.step())
    elapsed = time.perf_counter() - t0
    value = float(np.mean([r["sweep_cards_per_s"] for r in rates]))
    e2e = float(np.mean([r["e2e_cards_per_s"] for r in rates]))
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1000.0 * elapsed / max(1, args.steps),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": workload_config(args.cards, args.gpus, args.distinct, args.e2e_cards),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "physical_cores": phys, "kind": "port",
                         "sample": ref.sample_text(),
                         "angle_evals_per_s": float(np.mean([r["angle_evals_per_s"] for r in rates])),
                         "stage_seconds": {k: float(np.mean([r[k] for r in rates]))
                                           for k in ("t_decode_s", "t_pre_s", "t_angle_s", "t_warp_s", "t_encode_s")}},
        "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0,
                "what": "file to file: oracle imread -> correct_default -> imwrite(JPEG, 100)"},
        "note": "reference CPU path = C restatement of the reference's OpenCV call sequence (oracle/); the Rust "
                "crate + OpenCV 4.6 DLL cannot be built in this image",
    }
    print(json.dumps(line), flush=True)


def bind_to_gpu_numa(local_rank):
    """Runs this rank (and so its pinned host buffers, first-touch) on the NUMA node of its GPU."""
    try:
        import torch
        bus = torch.cuda.get_device_properties(local_rank).pci_bus_id
        dom = getattr(torch.cuda.get_device_properties(local_rank), "pci_domain_id", 0)
        dev_id = getattr(torch.cuda.get_device_properties(local_rank), "pci_device_id", 0)
        path = f"/sys/bus/pci/devices/{dom:04x}:{bus:02x}:{dev_id:02x}.0/numa_node"
        node = int(open(path).read().strip())
        if node < 0:
            return None
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            return node
    except Exception:
        pass
    return None


# --------------------------------------------------------------------------- GPU arm
class Timed:
    """CUDA-event timing of a region on `stream`, max over ranks."""

    def __init__(self, torch, dist, stream, dev, world):
        self.torch, self.dist, self.stream, self.dev, self.world = torch, dist, stream, dev, world

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def run(self, fn, steps):
        """-> seconds for `steps` calls of fn (max over ranks), last return value of fn."""
        torch = self.torch
        self.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(self.stream)
        out = None
        for _ in range(steps):
            out = fn()
        e1.record(self.stream)
        e1.synchronize()
        self.barrier()
        t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item()) / 1000.0, out


def kernel_rooflines(report, hbm_peak):
    """oics_profile_report rows -> per-kernel achieved GB/s against the measured HBM copy peak."""
    rows = []
    for r in report:
        ms, n, b = r["ms"], r["launches"], r["algorithmic_bytes"]
        row = {"kernel": r["kernel"], "launches": n, "ms_total": ms, "us_per_launch": 1000.0 * ms / max(1, n)}
        if b > 0 and ms > 0:
            row["algorithmic_bytes_per_launch"] = b / n
            row["achieved_gbs"] = b / (ms / 1000.0) / 1e9
            row["frac_of_hbm_peak"] = row["achieved_gbs"] / hbm_peak
        rows.append(row)
    return rows


def config5_image(ctx, api, synth):
    """The A4 generator scaled x4 with INTER_NEAREST: pre-processed at A4 on the GPU, replicated."""
    g, a = synth.base_card(5)
    bgr = ctx.rotate_mat(synth.to_bgr(g), -a, 1.0, api.INTER_LINEAR, api.WHITE, api.CLIP_DEFAULT)
    b, s = ctx.preprocess(bgr, api.PRE_OMR, 0, 0)
    big = np.repeat(np.repeat(b, 4, axis=0), 4, axis=1)
    big = np.ascontiguousarray(big[:14032, :9920])
    return np.ascontiguousarray(np.concatenate([big, big[:, -1:]], axis=1)), a      # 9921 x 14032


def run_gpu(args, rank, local_rank, world):
    import torch
    import torch.distributed as dist
    from omr_img_corrector_b200 import api, sharded, synth

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    numa = bind_to_gpu_numa(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    ctx = api.Context(local_rank)
    # the context's stream above the decoder's per-card streams (created at the default = lowest priority): when
    # an SM frees resources the sweep / encode CTAs of the main pipeline are placed first and the latency-bound
    # decode CTAs fill what is left (OICS_BENCH_STREAM_PRIORITY=0 restores equal priorities for A/B runs)
    prio = int(os.environ.get("OICS_BENCH_STREAM_PRIORITY", "-1"))
    stream = torch.cuda.Stream(device=dev, priority=prio)
    ctx.set_stream(stream.cuda_stream)
    timed = Timed(torch, dist, stream, dev, world)
    orc = None
    if rank == 0 and not args.no_oracle:
        from oracle import oracle as orc        # the checker: only ever used OUTSIDE the timed regions
    threads, phys = host_cores()

    # ---- synthetic cards: drawn on the host, skewed and pre-processed by the product kernels
    D = min(args.distinct, args.cards)
    N, angles, M = api.sweep_matrices(W, H, MAX_ANGLE, STEP, 1.0)
    skews, bgr_cards, bins = [], [], []
    for i in range(D):
        g, a = synth.base_card(rank * 100003 + i)
        bgr = ctx.rotate_mat(synth.to_bgr(g), -a, 1.0, api.INTER_LINEAR, api.WHITE, api.CLIP_DEFAULT)
        b, s = ctx.preprocess(bgr, api.PRE_OMR, 0, 0)
        skews.append(a)
        bins.append(b)
        if i < args.e2e_distinct:
            bgr_cards.append(bgr)
    cards = torch.empty((args.cards, H, W), dtype=torch.uint8, device=dev)
    for i in range(D):
        cards[i::D] = torch.from_numpy(bins[i]).to(dev)
    torch.cuda.synchronize()
    # pinned result buffers: the std-devs come back with an asynchronous copy, not through pageable staging
    out_v = torch.zeros((args.cards, N_ANGLES), dtype=torch.float64).pin_memory().numpy()
    out_h = torch.zeros((args.cards, N_ANGLES), dtype=torch.float64).pin_memory().numpy()

    def step():
        v, h = ctx.projection_sweep_batch(cards.data_ptr(), args.cards, W, H, M, out=(out_v, out_h))
        return api.rule_a_batch(h, v, N, STEP)     # exact rule A (omr.rs:147-221) per card, host

    for _ in range(max(3, args.warmup)):
        res = step()
    # sanity on every card: the sweep must recover the known synthetic skew (this is NOT the parity check)
    not_recovered = sum(1 for c in range(args.cards)
                        if res[1][c] != api.BELIEVED or abs(res[0][c] - skews[c % D]) > 0.1 + 1e-9)

    sampler = ClockSampler(local_rank)
    sampler.start()
    launches0 = ctx.launch_count
    sweep_ms, sweep_launches = [0.0], [0]

    def timed_step():
        step()
        ms, nl = ctx.last_sweep_kernel_ms()
        sweep_ms[0] += ms
        sweep_launches[0] += nl

    t_wall = time.perf_counter()
    elapsed_s, _ = timed.run(timed_step, args.steps)
    t_wall = time.perf_counter() - t_wall
    launches = ctx.launch_count - launches0
    clocks = sampler.stop()
    elapsed_ms = elapsed_s * 1000.0
    value = world * args.cards * args.steps / elapsed_s
    sweep_ms, sweep_launches = sweep_ms[0], sweep_launches[0]

    # ---- parity of the timed path against the oracle (rank 0, outside the timed region): every one of the 300
    # std-dev pairs of `oracle_cards` distinct cards bit for bit, and the rule-A angle / status
    parity = {"skew_recovered": {"cards_checked": args.cards, "not_recovered": not_recovered,
                                 "what": "sanity only: rule-A angle within 0.1 deg of the known synthetic skew and BELIEVED"}}
    if orc is not None:
        n_chk = min(args.oracle_cards, D)
        bad_sd = bad_angle = 0
        for i in range(n_chk):
            rv, rh = orc.projection_sweep(bins[i], M, threads=threads)
            ro = orc.rule_a(rh, rv, N, STEP)
            bad_sd += int(not (np.array_equal(out_v[i], rv) and np.array_equal(out_h[i], rh)))
            bad_angle += int((float(res[0][i]), int(res[1][i])) != (ro.angle, ro.status))
        parity["oracle"] = {"cards_checked": n_chk, "angles_per_card": N_ANGLES, "std_dev_mismatches": bad_sd,
                            "angle_or_status_mismatches": bad_angle,
                            "what": "v_sd/h_sd of all 300 candidates bit-identical to the oracle and the same rule-A result"}

    # ---- e2e: omr::correct_default through the C ABI on HOST buffers, copies inside the timed region
    E = args.e2e_cards
    e2e = None
    if E > 0:
        side_w, side_h = 3310, 4040          # CONTAIN size bound for |angle| <= 15 deg (3304 x 4031)
        host_out = torch.empty((E, side_w * side_h * 3), dtype=torch.uint8).pin_memory()
        outs = [host_out[i].numpy() for i in range(E)]
        ksteps = max(1, min(args.steps, args.e2e_steps))

        # (1) the headline: file to file (imread omr.rs:349 ... imwrite omr.rs:447-448), both codecs on the device
        distinct = [ctx.jpeg_encode(c, 100) for c in bgr_cards]
        fmax = max(len(d) for d in distinct)
        host_files = torch.empty((E, fmax), dtype=torch.uint8).pin_memory()
        files = []
        for i in range(E):
            d = distinct[i % len(distinct)]
            host_files[i, :len(d)] = torch.frombuffer(bytearray(d), dtype=torch.uint8)
            files.append(host_files[i].numpy()[:len(d)])

        # the reference runs correct_default on `MAX_WORKERS_COUNT` OS threads at once (thread_pool.rs:41-73); the
        # library's rule is one context per host thread (oics_cuda.h).  With --e2e-contexts T > 1 the files of a step
        # are split over T host threads, each with its own context and stream on this GPU, so one context's host
        # round trips (rule A between the sweep and the warp of a sub-batch) are covered by another's kernels.
        T = max(1, args.e2e_contexts)
        ctxs = [ctx] + [api.Context(local_rank) for _ in range(T - 1)]
        parts = [list(range(t, E, T)) for t in range(T)]

        def files_part(t):
            idx = parts[t]
            return ctxs[t].correct_batch_files([files[i] for i in idx], MAX_ANGLE, STEP, 0, 0, 125.0, 15.0,
                                               outs=[outs[i] for i in idx], jpeg_quality=100)

        def e2e_files_step():
            if T == 1:
                return files_part(0)
            res = [None] * T
            ths = [threading.Thread(target=lambda t=t: res.__setitem__(t, files_part(t))) for t in range(T)]
            [th.start() for th in ths]
            [th.join() for th in ths]
            for c in ctxs[1:]:
                c.synchronize()
            ang_, need_, stat_ = np.zeros(E), np.zeros(E, bool), np.zeros(E, np.int32)
            sizes_, fsz_ = np.zeros((E, 2), np.int32), np.zeros(E, np.int64)
            for t in range(T):
                a_, n_, s_, z_, f_ = res[t]
                ang_[parts[t]], need_[parts[t]], stat_[parts[t]], sizes_[parts[t]], fsz_[parts[t]] = a_, n_, s_, z_, f_
            return ang_, need_, stat_, sizes_, fsz_

        for _ in range(2):
            ang3, need3, stat3, sizes3, fsz3 = e2e_files_step()
        f_not_recovered = int(sum(1 for i in range(E) if stat3[i] != 0 or
                                  abs(ang3[i] - skews[i % len(bgr_cards)]) > 0.1 + 1e-9))
        oracle_files = None
        if orc is not None:       # one file, whole pipeline, byte for byte (oracle sweep on all host threads)
            src = orc.jpeg_decode(distinct[0])
            oa, oneed, oimg = orc.correct_default(src, MAX_ANGLE, STEP, 0, 0, 125.0, 15.0, threads=threads)
            ofile = orc.jpeg_encode(oimg, 100)
            oracle_files = {"files_checked": 1,
                            "angle_equal": bool(oa == float(ang3[0]) and bool(oneed) == bool(need3[0])),
                            "output_file_bytes_equal": bool(outs[0][:int(fsz3[0])].tobytes() == ofile)}
        t_k, _ = timed.run(e2e_files_step, ksteps)
        e2e = {"value": world * E * ksteps / t_k, "unit": UNIT,
               "what": "oics_correct_batch_files: JPEG (quality 100) files in pinned host memory in, JPEG files out; "
                       "decode, deskew and encode on the device",
               "h2d_bytes_per_step": int(sum(f.size for f in files)), "d2h_bytes_per_step": int(fsz3.sum()) + E * 16,
               "cards_per_gpu_per_step": E, "steps": ksteps, "host_threads_and_contexts_per_gpu": T,
               "skew_not_recovered": f_not_recovered,
               "oracle": oracle_files, "numa_node": numa}

        # (2) side fields: raw pixels over PCIe (26 MB in + 33 MB out per card), fewer steps
        host_in = torch.empty((E, H, W, 3), dtype=torch.uint8).pin_memory()
        hin = host_in.numpy()
        for i in range(E):
            hin[i] = bgr_cards[i % len(bgr_cards)]
        ssteps = max(1, min(ksteps, 3))

        def e2e_step():
            return ctx.correct_batch(hin, MAX_ANGLE, STEP, 0, 0, 125.0, 15.0, outs=outs)

        for _ in range(2):
            ang, need, stat, sizes = e2e_step()
        raw_bad = int(sum(1 for i in range(E) if stat[i] != 0 or ang[i] != ang3[i]))
        t_e2e, _ = timed.run(e2e_step, ssteps)
        d2h = int(sum(int(sizes[i][0]) * int(sizes[i][1]) * 3 for i in range(E)))
        e2e["raw_pixels"] = {"value": world * E * ssteps / t_e2e, "unit": UNIT,
                             "what": "oics_correct_batch: pinned BGR cards in, deskewed BGR cards out",
                             "h2d_bytes_per_step": int(E * W * H * 3), "d2h_bytes_per_step": d2h + E * 16,
                             "steps": ssteps, "angle_differs_from_file_path": raw_bad}

        def e2e_jpeg_step():
            return ctx.correct_batch(hin, MAX_ANGLE, STEP, 0, 0, 125.0, 15.0, outs=outs, jpeg_quality=100)

        for _ in range(2):
            ang2, need2, stat2, sizes2, fsz = e2e_jpeg_step()
        jbad = int(sum(1 for i in range(E) if stat2[i] != 0 or ang2[i] != ang[i]))
        t_j, _ = timed.run(e2e_jpeg_step, ssteps)
        e2e["jpeg_out"] = {"value": world * E * ssteps / t_j, "unit": UNIT,
                           "what": "oics_correct_batch_jpeg: BGR pixels in, the JPEG (quality 100) file of every deskewed card out",
                           "h2d_bytes_per_step": int(E * W * H * 3), "d2h_bytes_per_step": int(fsz.sum()) + E * 16,
                           "steps": ssteps, "angle_differs_from_pixel_path": jbad}
        del host_in, hin

    # ---- per-kernel device times of the e2e path and of the headline step, from the library's own CUDA events
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        hbm_peak, hbm_src = float(json.load(open(peaks_path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    else:
        hbm_peak, hbm_src = 6650.0, "fallback (B200_PROFILING.md)"
    roofline_e2e = None
    if E > 0 and rank == 0:
        ctx.profile(True)
        e2e_files_step()
        roofline_e2e = {"peak_gbs": hbm_peak, "peak_source": hbm_src,
                        "source": "oics_profile: CUDA events around every launch of one untimed oics_correct_batch_files "
                                  f"step ({E} cards) after the timed region; algorithmic bytes per DESIGN.md section 4",
                        "kernels": kernel_rooflines(ctx.profile_report(), hbm_peak)}
        ctx.profile(False)

    # ---- the other BASELINE configs, driver-visible (untimed by the headline; each has its own events)
    configs = {}
    if not args.no_configs:
        configs = run_configs(args, ctx, api, sharded, synth, orc, timed, torch, dist, dev, rank, world, threads, hbm_peak)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel (the fused sweep): 1 algorithmic byte per (angle, pixel) sample,
    # against the MEASURED L2 read peak (the sweep re-reads one card 2N times, its ceiling is on-chip)
    bytes_per_launch = float(args.cards) * args.steps * N_ANGLES * W * H / max(1, sweep_launches)
    avg_launch_s = (sweep_ms / 1000.0) / max(1, sweep_launches)
    achieved = bytes_per_launch / avg_launch_s / 1e9
    cards_per_launch = args.cards * args.steps / max(1, sweep_launches)
    try:
        l2_gbs = ctx.probe_l2_read(32 << 20, 64)
        smem_gbs = ctx.probe_smem_read(4096)
    except Exception as exc:  # a probe failure must not lose the bench line
        l2_gbs = smem_gbs = None
        probe_err = str(exc)
    traffic, traffic_note = None, "no ncu capture committed for this launch shape"
    tp = os.path.join(ROOT, "profiles", "sweep_traffic.json")
    if os.path.exists(tp):
        try:   # ncu --set full capture of one launch: dram__bytes_read.sum + dram__bytes_write.sum
            tj = json.load(open(tp))
            per_card = float(tj["dram_bytes_per_launch"]) / float(tj["cards_per_launch"])
            if abs(float(tj["cards_per_launch"]) - cards_per_launch) < 0.5:
                traffic, traffic_note = float(tj["dram_bytes_per_launch"]), f"ncu capture at this launch shape ({tj.get('capture')})"
            else:
                traffic = per_card * cards_per_launch
                traffic_note = (f"ncu capture at {tj['cards_per_launch']} cards per launch scaled per card to "
                                f"{cards_per_launch:.1f} ({tj.get('capture')})")
        except Exception:
            traffic = None
    roofline = {"bound": "l2", "achieved": achieved, "peak": l2_gbs, "unit": "GB/s",
                "frac": (achieved / l2_gbs) if l2_gbs else None, "traffic": traffic, "traffic_note": traffic_note,
                "kernel": "oics::sweepw_kernel",
                "peak_source": "measured live after the timed region: oics_probe_l2_read (32 MB L2-resident buffer, 64 passes, "
                               "ld.global.cg.v4, every SM)" if l2_gbs else f"probe failed: {probe_err}",
                "algorithmic_bytes_per_launch": bytes_per_launch, "cards_per_launch": cards_per_launch,
                "avg_launch_ms": avg_launch_s * 1000.0, "launches": sweep_launches,
                "kernel_share_of_step": sweep_ms / elapsed_ms,
                "hbm": {"peak": hbm_peak, "peak_source": hbm_src,
                        "achieved_dram_gbs": (traffic / avg_launch_s / 1e9) if traffic else None,
                        "frac": (traffic / avg_launch_s / 1e9 / hbm_peak) if traffic else None,
                        "algorithmic_over_hbm_peak": achieved / hbm_peak,
                        "note": "real DRAM traffic of a launch over the measured copy peak: the kernel is not HBM-limited"},
                "smem": {"peak": smem_gbs, "frac": (achieved / smem_gbs) if smem_gbs else None,
                         "source": "oics_probe_smem_read (64 KB tile per CTA, LDS.128)"},
                "note": "1 B per (angle, pixel) sample = the rotated byte image the reference writes and re-reads "
                        "(2N*w*h per card, SURVEY 8(d)); the kernel assembles 32 samples per lane-word from the 1.1 MB "
                        "bit plane of the card staged in shared memory, so the L2 figure is the accounting ceiling "
                        "SURVEY 8(d) fixes, and the physical limiters are issue slots and the shared-memory crossbar (ncu)"}

    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        ref = CpuReference(threads)
        r = ref.step()
        cpu = {"value": r["sweep_cards_per_s"], "unit": UNIT, "cores": threads, "physical_cores": phys, "kind": "port",
               "sample": ref.sample_text(), "e2e_value": r["e2e_cards_per_s"],
               "angle_evals_per_s": r["angle_evals_per_s"], "cpu_seconds": r["cpu_seconds"],
               "stage_seconds": {k: r[k] for k in ("t_decode_s", "t_pre_s", "t_angle_s", "t_warp_s", "t_encode_s")}}

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": max(3, args.warmup), "ms_per_step": elapsed_ms / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": workload_config(args.cards, world, D, E),
        "angle_evals_per_s": value * N_ANGLES,
        "samples_per_s": value * N_ANGLES * W * H,
        "roofline": roofline, "roofline_e2e": roofline_e2e, "cpu_baseline": cpu, "e2e": e2e,
        "gpu_launches": int(launches), "clocks": clocks, "wall_s": t_wall,
        "parity": parity, "configs": configs,
    }
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def run_configs(args, ctx, api, sharded, synth, orc, timed, torch, dist, dev, rank, world, threads, hbm_peak):
    """BASELINE configs 2, 3 (rank 0, one A4 card each), 5 (all ranks) and the reference's default operating point
    (rank 0).  Every number is device-timed with CUDA events; every oracle comparison runs outside the timed call."""
    out = {}
    stream = timed.stream

    def ev_time(fn, reps=1):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        r = None
        for _ in range(reps):
            r = fn()
        e1.record(stream)
        e1.synchronize()
        return e0.elapsed_time(e1) / 1000.0 / reps, r

    if rank == 0:
        g, a = synth.base_card(0)
        bgr = ctx.rotate_mat(synth.to_bgr(g), -a, 1.0, api.INTER_LINEAR, api.WHITE, api.CLIP_DEFAULT)
        d_bgr = torch.from_numpy(bgr).to(dev)
        # config 2: Hough-line deskew of the card (omr.rs:231-301): Canny on BGR, HoughLinesP(thr 0, 125/15), mode
        ctx.result_from_edges_detection(d_bgr, 125.0, 15.0)
        ctx.profile(True)
        t2, r2 = ev_time(lambda: ctx.result_from_edges_detection(d_bgr, 125.0, 15.0))
        k2 = kernel_rooflines(ctx.profile_report(), hbm_peak)
        ctx.profile(False)
        c2 = {"seconds": t2, "angle": r2.angle, "status": r2.status, "skew": a, "kernels": k2,
              "what": "omr::get_result_from_edges_detection on the device-resident A4 BGR card (Canny + HoughLinesP + angle mode)"}
        if orc is not None:
            o2 = orc.get_result_from_edges_detection(bgr, 125.0, 15.0)
            c2["oracle_equal"] = bool((o2.angle, o2.status, o2.candidates) == (r2.angle, r2.status, r2.candidates))
        # the same with 32 cards in flight (one thread-block cluster per card in one PPHT launch): the reference's
        # thread-pool concurrency (thread_pool.rs:41-73) for this estimator
        nb2 = args.hough_batch
        if nb2 > 0:
            d32 = d_bgr[None].repeat(nb2, 1, 1, 1).contiguous()
            ctx.result_from_edges_detection_batch(d32, 125.0, 15.0)
            tb, rb = ev_time(lambda: ctx.result_from_edges_detection_batch(d32, 125.0, 15.0))
            c2["batch"] = {"cards_in_flight": nb2, "seconds": tb, "cards_per_s": nb2 / tb,
                           "equal_to_single_call": bool(all(float(x) == r2.angle for x in rb[0]) and
                                                        all(int(x) == r2.status for x in rb[1]) and not rb[2].any())}
            del d32
        out["c2_hough"] = c2
        # config 3: FFT dominant-line deskew (fft.rs:145-256) of the gray card
        d_gray = torch.from_numpy(np.ascontiguousarray(bgr[:, :, 0])).to(dev)     # the card is gray-in-BGR
        ctx.angle_with_fft(d_gray, 125.0, 150.0, 125.0, 15.0)
        ctx.profile(True)
        t3, r3 = ev_time(lambda: ctx.angle_with_fft(d_gray, 125.0, 150.0, 125.0, 15.0))
        k3 = kernel_rooflines(ctx.profile_report(), hbm_peak)
        ctx.profile(False)
        c3 = {"seconds": t3, "angle": r3, "kernels": k3,
              "what": "fft::get_angle_with_fft on the device-resident gray A4 card (FFT image + Canny + HoughLinesP(100) + vote)"}
        if orc is not None:
            c3["oracle_equal"] = bool(orc.get_angle_with_fft(np.ascontiguousarray(bgr[:, :, 0]), 125.0, 150.0, 125.0, 15.0) == r3)
        out["c3_fft"] = c3
        del d_bgr, d_gray

        # the reference's shipped operating point (lib.rs:92-97): 1240x1150 cards, +-45 deg at 0.2 deg, shrink to 248x230
        # skews spread over the test's own range (lib.rs:151-152: -45 .. 45 deg), so the Hough fallback of
        # omr.rs:366-403 is exercised at its natural rate
        P = DEFAULT_OP
        nd, ncards = 32, args.default_op_cards
        small, sk = [], []
        rng = np.random.default_rng(20261105)
        bases = []
        for i in range(8):
            g, _ = synth.base_card(1000 + i)
            bases.append(synth.to_bgr(np.ascontiguousarray(g[200:1350, 100:1340])))
        for i in range(nd):
            a_i = round(float(rng.uniform(-44.0, 44.0)), 1)
            small.append(ctx.rotate_mat(bases[i % 8], -a_i, 1.0, api.INTER_LINEAR, api.WHITE, api.CLIP_DEFAULT))
            sk.append(a_i)
        h_in = torch.empty((ncards, 1150, 1240, 3), dtype=torch.uint8).pin_memory()
        for i in range(ncards):
            h_in[i] = torch.from_numpy(small[i % nd])
        side = int(np.ceil((1240 + 1150) * 0.70710678118654757)) + 2
        h_out = torch.empty((ncards, side * side * 3), dtype=torch.uint8).pin_memory()
        h_outs = [h_out[i].numpy() for i in range(ncards)]
        d_in = h_in.to(dev)
        d_out = torch.empty((ncards, side * side * 3), dtype=torch.uint8, device=dev)
        d_outs = [d_out[i] for i in range(ncards)]
        host_fn = lambda: ctx.correct_batch(h_in.numpy(), *P, outs=h_outs)
        dev_fn = lambda: ctx.correct_batch(d_in, *P, outs=d_outs)
        host_fn(); dev_fn()
        t_h, rh_ = ev_time(host_fn, 3)
        t_d, rd_ = ev_time(dev_fn, 3)
        dop = {"params": list(P), "cards": ncards, "card_size": [1240, 1150],
               "host_cards_per_s": ncards / t_h, "device_resident_cards_per_s": ncards / t_d,
               "need_check": int(np.sum(rd_[1])), "errors": int(np.sum(rd_[2] != 0)),
               "within_0.4_deg_of_skew": int(sum(1 for i in range(ncards) if abs(float(rd_[0][i]) - sk[i % nd]) <= 0.4 + 1e-9)),
               "skews": "32 distinct cards, skews uniform in [-44, 44] deg (the reference test sweeps -45 .. 45)",
               "what": "oics_correct_batch at the reference's default parameters; host = pinned BGR in / deskewed BGR out, "
                       "device_resident = cards and outputs in HBM"}
        if orc is not None:
            eq = 0
            chk = list(range(0, nd, 4))
            for i in chk:
                oa, oneed, oimg = orc.correct_default(small[i], *P)
                dw, dh = int(rh_[3][i][0]), int(rh_[3][i][1])
                same = (oa == float(rh_[0][i]) and bool(oneed) == bool(rh_[1][i]) and oimg.shape == (dh, dw, 3)
                        and np.array_equal(h_outs[i][:dw * dh * 3].reshape(dh, dw, 3), oimg)
                        and float(rd_[0][i]) == oa)
                eq += int(same)
            dop["oracle_equal"] = eq
            dop["oracle_checked"] = len(chk)
        out["default_op"] = dop
        del h_in, h_out, d_in, d_out

    # config 5: every rank holds the scan, evaluates its slice of the 9000 angles (oics_sweep_sharded writes into
    # the device exchange buffer), ONE NCCL all-gather, exact rule A on every rank
    if not args.no_config5:
        big, a5 = config5_image(ctx, api, synth)
        d_big = torch.from_numpy(big).to(dev)
        hh, ww = big.shape
        N5, angles5, M5 = api.sweep_matrices(ww, hh, 45, 0.01, 1.0)
        per = api.shard_layout(len(M5), world)
        sd = torch.zeros((world, 2, per), dtype=torch.float64, device=dev)

        def c5():
            v, h = sharded.sweep_sharded_device(ctx, d_big, M5, rank, world, None, dev, sd)
            return v, h, api.rule_a(h, v, N5, 0.01)

        c5()                      # warm-up: tables, NCCL
        t5, (v5, h5, r5) = timed.run(c5, 1)
        c5r = {"seconds": t5, "n_gpus": world, "angles": len(M5), "image": [ww, hh],
               "samples_per_s": len(M5) * float(ww) * hh / t5, "angle": r5.angle, "status": r5.status, "skew": a5,
               "what": "one 9921x14032 scan, +-45 deg at 0.01 deg: oics_sweep_sharded on every rank, one NCCL all-gather "
                       "of the exchange buffer (16 B per angle), exact rule A on every rank"}
        if rank == 0:
            if world > 1:
                t1, (v1, h1) = ev_time(lambda: ctx.projection_sweep_batch(d_big.data_ptr(), 1, ww, hh, M5))
                r1 = api.rule_a(h1[0], v1[0], N5, 0.01)
                c5r["single_gpu_seconds"] = t1
                c5r["identical_to_single_gpu"] = bool(np.array_equal(v1[0], v5) and np.array_equal(h1[0], h5) and
                                                     (r1.angle, r1.status) == (r5.angle, r5.status))
            if orc is not None:
                idx = [0, 2500, 4500 + int(round(a5 * 100)), 4501, 6230, 8999]
                rv, rh = orc.projection_sweep(big, M5[idx], threads=min(threads, len(idx)))
                c5r["oracle_angles_checked"] = len(idx)
                c5r["oracle_equal"] = bool(np.array_equal(v5[idx], rv) and np.array_equal(h5[idx], rh))
        if world > 1:
            dist.barrier()
        out["c5_huge_scan"] = c5r
        del d_big, big
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--cards", type=int, default=512, help="binary cards per GPU per step (4096 / 8)")
    ap.add_argument("--distinct", type=int, default=16, help="distinct synthetic cards per GPU (cycled)")
    ap.add_argument("--e2e-cards", dest="e2e_cards", type=int, default=128)
    ap.add_argument("--e2e-distinct", dest="e2e_distinct", type=int, default=8)
    ap.add_argument("--e2e-steps", dest="e2e_steps", type=int, default=10)
    ap.add_argument("--e2e-contexts", dest="e2e_contexts", type=int, default=1,
                    help="host threads (one context each) that share a GPU in the file-to-file e2e")
    ap.add_argument("--oracle-cards", dest="oracle_cards", type=int, default=2,
                    help="distinct cards whose 300 std-dev pairs are compared with the oracle (rank 0, untimed)")
    ap.add_argument("--default-op-cards", dest="default_op_cards", type=int, default=128)
    ap.add_argument("--hough-batch", dest="hough_batch", type=int, default=36, help="A4 cards in flight in the config-2 batch line (0 = skip); 36 clusters of 4 CTAs fill 144 of the 148 SMs")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="skip the configs 2 / 3 / 5 / default-op block")
    ap.add_argument("--no-config5", action="store_true")
    ap.add_argument("--no-oracle", action="store_true", help="skip every oracle comparison (profiling runs)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    run_gpu(args, rank, local_rank, world)


if __name__ == "__main__":
    main()
