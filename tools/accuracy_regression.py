#!/usr/bin/env python
"""the reference's own acceptance run (packages/lib/src/lib.rs:131-245)
on the GPU path, at the reference's operating point.

For every card and every skew a in [-45, 45) the input is synthesised exactly like the test
does (rotate by -a, INTER_LINEAR, white border, DEFAULT clip; gray -> GRAY2RGB is the identity
on these gray-replicated cards), then omr::correct_default runs with the test's parameter
tuple (45, 0.2, 248, 230, 150.0, 50.0) through oics_correct_batch, and the result is
classified with the test's own vocabulary (lib.rs:214-226):
    need_check -> NOT_BELIEVED, |a - result| > 0.5 -> ERROR, > 0.4 -> NOT_SO_RIGHT, else SUCCESS.
A sub-sample of the inputs is re-run through the CPU oracle and (angle, need_check) must be
identical.  The reference's dataset JPEGs do not travel to the GPU box, so the cards are the
synthetic 1240x1150 "dataset-like" crops the parity tests use.  The test writes every skewed card
as ./tmp.jpg with quality 100 and hands that FILE to correct_default (lib.rs:185-192), which writes
the deskewed card as a JPEG again: with --files (default) the run does the same -- the card is
encoded on the device (oics_jpeg_encode), the files go through oics_correct_batch_files (decode,
deskew, encode on the device), and the oracle check is oracle-imread -> oracle correct_default ->
oracle imwrite, compared byte for byte.  --no-files runs the pixel path (oics_correct_batch).

    python tools/accuracy_regression.py [--cards 2] [--skew-step 0.1] [--oracle-every 45]
Prints one JSON object.
"""
import argparse
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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cards", type=int, default=2)
    ap.add_argument("--skew-step", type=float, default=0.1)
    ap.add_argument("--oracle-every", type=int, default=45, help="check every k-th input against the CPU oracle (0 = none)")
    ap.add_argument("--chunk", type=int, default=64)
    ap.add_argument("--files", dest="files", action="store_true", default=True)
    ap.add_argument("--no-files", dest="files", action="store_false")
    args = ap.parse_args()

    ctx = api.Context(0)
    n_skews = int(round(90.0 / args.skew_step))
    skews = [(-n_skews // 2 + i) * args.skew_step for i in range(n_skews)]     # test_iter_idx * 0.1, lib.rs:151-152
    hist = {"SUCCESS": 0, "NOT_SO_RIGHT": 0, "ERROR": 0, "NOT_BELIEVED": 0, "error_status": 0}
    total_mistake, n_total, t_gpu = 0.0, 0, 0.0
    mismatches, n_oracle, t_cpu = [], 0, 0.0
    orc = None
    if args.oracle_every > 0:
        from oracle import oracle as orc
    for c in range(args.cards):
        g, _ = synth.base_card(100 + c)
        base = synth.to_bgr(np.ascontiguousarray(g[200:1350, 100:1340]))      # 1240 x 1150, like the dataset images
        for k0 in range(0, n_skews, args.chunk):
            sk = skews[k0:k0 + args.chunk]
            cards = np.stack([ctx.rotate_mat(base, -a, 1.0, api.INTER_LINEAR, api.WHITE, api.CLIP_DEFAULT) for a in sk])
            files = outs = fsz = None
            if args.files:
                files = [np.frombuffer(ctx.jpeg_encode(cards[i], 100), np.uint8) for i in range(len(sk))]   # ./tmp.jpg
                outs = [np.empty(3 << 20, np.uint8) for _ in sk]
            t0 = time.perf_counter()
            if args.files:
                ang, need, stat, _, fsz = ctx.correct_batch_files(files, *PARAMS, outs=outs, jpeg_quality=100)
            else:
                ang, need, stat, _ = ctx.correct_batch(cards, *PARAMS)
            t_gpu += time.perf_counter() - t0
            for i, a in enumerate(sk):
                n_total += 1
                if stat[i] != 0:
                    hist["error_status"] += 1
                    continue
                hist[classify(a, float(ang[i]), bool(need[i]))] += 1
                total_mistake += abs(a - float(ang[i]))
                if orc is not None and (k0 + i) % args.oracle_every == 0:
                    t0 = time.perf_counter()
                    if args.files:
                        src = orc.jpeg_encode(cards[i], 100)
                        ra, rneed, rout = orc.correct_default(orc.jpeg_decode(src), *PARAMS)
                        same_files = src == files[i].tobytes() and orc.jpeg_encode(rout, 100) == outs[i][:int(fsz[i])].tobytes()
                    else:
                        ra, rneed, _ = orc.correct_default(cards[i], *PARAMS)
                        same_files = True
                    t_cpu += time.perf_counter() - t0
                    n_oracle += 1
                    if (ra, bool(rneed)) != (float(ang[i]), bool(need[i])) or not same_files:
                        mismatches.append({"card": c, "skew": a, "gpu": [float(ang[i]), bool(need[i])], "oracle": [ra, bool(rneed)]})
    out = {
        "what": "lib.rs:131-245 acceptance run on the CUDA path (synthetic dataset-like cards), "
                + ("file to file like the test: tmp.jpg (q 100) in, deskewed JPEG out, codecs on the device" if args.files
                   else "pixel path, no JPEG round trip"),
        "params": PARAMS, "cards": args.cards, "skews_per_card": n_skews, "skew_step": args.skew_step, "inputs": n_total,
        "histogram": hist, "average_error_deg": total_mistake / max(1, n_total - hist["error_status"]),
        "gpu_seconds_correct_batch": t_gpu, "gpu_cards_per_s": n_total / t_gpu if t_gpu else None,
        "oracle_checked": n_oracle, "oracle_mismatches": mismatches,
        "oracle_seconds_per_card": t_cpu / n_oracle if n_oracle else None,
    }
    print(json.dumps(out, indent=1))
    return 1 if mismatches else 0


if __name__ == "__main__":
    sys.exit(main())
