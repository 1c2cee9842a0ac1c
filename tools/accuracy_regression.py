#!/usr/bin/env python
"""the reference's own acceptance run
(`crate_omr_correct_default_all_situation_test`, packages/lib/src/lib.rs:131-245) on the GPU path,
at the reference's operating point, on the reference's own fixtures.

For every card of the dataset (tests/fixtures/dataset/*.jpg = the reference's dataset/dataset,
lib.rs:143-150) and every skew a = k * 0.1, k in -450..450 (lib.rs:151-152) the input is synthesised
exactly like the test does: imread(IMREAD_COLOR), rotate by -a (INTER_LINEAR, white border, DEFAULT clip,
lib.rs:154-166), gray -> GRAY2RGB (the identity on these gray-replicated cards, lib.rs:168-186), written as
./tmp.jpg with quality 100 (lib.rs:188-190).  That FILE goes through omr::correct_default with the test's
parameter tuple (45, 0.2, 248, 230, 150.0, 50.0) (lib.rs:194-205) -- here oics_correct_batch_files: decode,
deskew and re-encode on the device -- and the result is classified with the test's own vocabulary
(lib.rs:214-226): need_check -> NOT_BELIEVED, |a - result| > 0.5 -> ERROR, > 0.4 -> NOT_SO_RIGHT, else SUCCESS.
Every `--oracle-every`-th input (>= 1 % by default) is re-run through the CPU oracle -- oracle imread ->
correct_default -> imwrite -- and (angle, need_check, input file bytes, output file bytes) must be identical.

    python tools/accuracy_regression.py [--dataset tests/fixtures/dataset] [--cards N] [--skew-step 0.1]
Without --dataset (or when the directory is empty) the cards are the synthetic 1240x1150 "dataset-like" crops
the parity tests use.  --no-files runs the pixel path (oics_correct_batch).  Prints one JSON object.
"""
import argparse
import concurrent.futures as cf
import glob
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from omr_img_corrector_b200 import api, synth  # noqa: E402

PARAMS = (45, 0.2, 248, 230, 150.0, 50.0)     # lib.rs:92-97 / 196-203


def classify(skew, angle, need):
    d = abs(skew - angle)
    if need:
        return "NOT_BELIEVED"
    if d > 0.5:
        return "ERROR"
    if d > 0.4:
        return "NOT_SO_RIGHT"
    return "SUCCESS"


def load_cards(ctx, dataset, n_cards):
    """-> list of (name, BGR card).  Dataset files are read with the device decoder (imread IMREAD_COLOR)."""
    files = sorted(glob.glob(os.path.join(dataset, "*.jpg"))) if dataset else []
    if files:
        if n_cards > 0:
            files = files[:: max(1, len(files) // n_cards)][:n_cards]
        out = []
        for f in files:
            raw = np.fromfile(f, np.uint8)
            out.append((os.path.basename(f), np.array(ctx.jpeg_decode(raw), copy=True), raw.tobytes()))
        return out, "reference dataset"
    out = []
    for c in range(max(1, n_cards)):
        g, _ = synth.base_card(100 + c)
        out.append((f"synthetic{c}", synth.to_bgr(np.ascontiguousarray(g[200:1350, 100:1340])), None))
    return out, "synthetic dataset-like cards (fixture directory absent)"


def run(ctx, cards, skews, files_mode=True, oracle_every=97, chunk=450, oracle_threads=None, progress=False):
    import torch
    hist = {"SUCCESS": 0, "NOT_SO_RIGHT": 0, "ERROR": 0, "NOT_BELIEVED": 0, "error_status": 0}
    total_mistake, n_total, t_gpu = 0.0, 0, 0.0
    checks = []          # (card idx, skew, input file bytes, output file bytes, gpu angle, gpu need)
    dev = torch.device("cuda", ctx.device)
    counter = 0
    for ci, (name, base, _raw) in enumerate(cards):
        h, w = base.shape[:2]
        d_base = torch.from_numpy(np.ascontiguousarray(base)).to(dev)
        d_rot = torch.empty((chunk, h, w, 3), dtype=torch.uint8, device=dev)
        side = int(np.ceil((w + h) * 0.70710678118654757)) + 2
        pin_files = torch.empty((chunk, 4096 + 2 * w * h), dtype=torch.uint8).pin_memory()
        pin_outs = torch.empty((chunk, side * side * 3 // 2 + 4096 if files_mode else side * side * 3), dtype=torch.uint8).pin_memory()
        for k0 in range(0, len(skews), chunk):
            sk = skews[k0:k0 + chunk]
            files = []
            for i, a in enumerate(sk):        # lib.rs:154-190: rotate by -a, INTER_LINEAR, white, DEFAULT clip; tmp.jpg q100
                M, dw, dh = api.rotate_geometry(w, h, -a, 1.0, api.CLIP_DEFAULT)
                ctx.warp_affine(d_base, M, (w, h), api.INTER_LINEAR, api.WHITE, out=d_rot[i])
                if files_mode:
                    data = ctx.jpeg_encode(d_rot[i], 100, out=pin_files[i].numpy())
                    files.append(pin_files[i].numpy()[:len(data)])
            outs = [pin_outs[i].numpy() for i in range(len(sk))]
            t0 = time.perf_counter()
            if files_mode:
                ang, need, stat, _, fsz = ctx.correct_batch_files(files, *PARAMS, outs=outs, jpeg_quality=100)
            else:
                ang, need, stat, sizes = ctx.correct_batch(d_rot[:len(sk)], *PARAMS, outs=outs)
            t_gpu += time.perf_counter() - t0
            for i, a in enumerate(sk):
                n_total += 1
                if stat[i] != 0:
                    hist["error_status"] += 1
                    continue
                hist[classify(a, float(ang[i]), bool(need[i]))] += 1
                total_mistake += abs(a - float(ang[i]))
                if oracle_every > 0 and counter % oracle_every == 0:
                    if files_mode:
                        checks.append((ci, a, files[i].tobytes(), outs[i][:int(fsz[i])].tobytes(), float(ang[i]), bool(need[i])))
                    else:
                        checks.append((ci, a, d_rot[i].cpu().numpy(), None, float(ang[i]), bool(need[i])))
                counter += 1
        if progress:
            print(f"card {ci + 1}/{len(cards)} {name}: {n_total} inputs, {t_gpu:.1f} s in correct_batch", file=sys.stderr, flush=True)
    mismatches, t_cpu = [], 0.0
    if checks:
        from oracle import oracle as orc
        orc.lib()

        def check(item):
            ci, a, src, out, gang, gneed = item
            t0 = time.perf_counter()
            # the whole chain again on the CPU, from the card: rotate (lib.rs:154-166) -> tmp.jpg -> correct_default -> output file
            skewed = orc.rotate_mat(cards[ci][1], -a, 1.0, 1, orc.WHITE, 0)
            if files_mode:
                tmp = orc.jpeg_encode(skewed, 100)
                ra, rneed, rout = orc.correct_default(orc.jpeg_decode(tmp), *PARAMS)
                same = tmp == src and orc.jpeg_encode(rout, 100) == out
            else:
                ra, rneed, rout = orc.correct_default(skewed, *PARAMS)
                same = bool(np.array_equal(skewed, src))
            ok = (ra, bool(rneed)) == (gang, gneed) and same
            return ok, {"card": cards[ci][0], "skew": a, "gpu": [gang, gneed], "oracle": [ra, bool(rneed)], "files_equal": bool(same)}, time.perf_counter() - t0

        with cf.ThreadPoolExecutor(max_workers=oracle_threads or len(os.sched_getaffinity(0))) as ex:
            for ok, info, dt in ex.map(check, checks):
                t_cpu += dt
                if not ok:
                    mismatches.append(info)
            # imread of the dataset file itself (device decoder) against the oracle's, for every card that was sampled
            def check_decode(ci):
                name, base, raw = cards[ci]
                return raw is None or bool(np.array_equal(orc.jpeg_decode(raw), base)), name
            for ok, name in ex.map(check_decode, sorted({c[0] for c in checks})):
                if not ok:
                    mismatches.append({"card": name, "imread_differs": True})
    return {"histogram": hist, "inputs": n_total, "average_error_deg": total_mistake / max(1, n_total - hist["error_status"]),
            "gpu_seconds_correct_batch": t_gpu, "gpu_cards_per_s": n_total / t_gpu if t_gpu else None,
            "oracle_checked": len(checks), "oracle_mismatches": mismatches,
            "oracle_seconds_per_card": t_cpu / len(checks) if checks else None}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--dataset", default=os.path.join(ROOT, "tests", "fixtures", "dataset"))
    ap.add_argument("--cards", type=int, default=0, help="number of cards (0 = every dataset card; 2 synthetic ones without a dataset)")
    ap.add_argument("--skew-step", type=float, default=0.1)
    ap.add_argument("--oracle-every", type=int, default=97, help="check every k-th input against the CPU oracle (0 = none)")
    ap.add_argument("--chunk", type=int, default=450, help="inputs per oics_correct_batch_files call (the fallback queue is flushed per call)")
    ap.add_argument("--files", dest="files", action="store_true", default=True)
    ap.add_argument("--no-files", dest="files", action="store_false")
    args = ap.parse_args()

    ctx = api.Context(0)
    cards, source = load_cards(ctx, args.dataset, args.cards if args.cards else (0 if glob.glob(os.path.join(args.dataset, "*.jpg")) else 2))
    n_skews = int(round(90.0 / args.skew_step))
    skews = [(-n_skews // 2 + i) * args.skew_step for i in range(n_skews)]     # test_iter_idx * 0.1, lib.rs:151-152
    res = run(ctx, cards, skews, args.files, args.oracle_every, args.chunk, progress=True)
    out = {"what": "lib.rs:131-245 acceptance run on the CUDA path, "
                   + ("file to file like the test: tmp.jpg (q 100) in, deskewed JPEG out, codecs on the device" if args.files
                      else "pixel path, no JPEG round trip"),
           "cards_source": source, "params": PARAMS, "cards": len(cards), "skews_per_card": n_skews, "skew_step": args.skew_step}
    out.update(res)
    print(json.dumps(out, indent=1))
    return 1 if res["oracle_mismatches"] else 0


if __name__ == "__main__":
    sys.exit(main())
