#!/usr/bin/env python
"""Per-kernel roofline table of configs 1-3 on one A4 card, from the ncu launch list that
tools/gpu_kernels.sh writes (gpu__time_duration + DRAM bytes per launch, cold cache, serialised).

Algorithmic bytes per launch follow SURVEY.md section 8(d) / DESIGN.md section 4; the peak is
MEASURED_PEAKS.json `hbm_gbs` (fallback 6650 GB/s, B200_PROFILING.md).

    python tools/kernel_rooflines.py gpurun_out/launches_configs.csv [out.json]
"""
import collections
import csv
import json
import math
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
W, H = 2480, 3508
SKEW = 9.1   # CARD=1 of tools/configs_bench.py


def contain_size(w, h, angle):
    r = angle * math.pi / 180.0
    return (math.ceil(h * abs(math.sin(r)) + w * abs(math.cos(r))), math.ceil(w * abs(math.sin(r)) + h * abs(math.cos(r))))


def algorithmic_bytes():
    px = W * H
    cw, chh = contain_size(W, H, SKEW)
    return {
        "gray_erode_bits_kernel<3>": (3 * px + px, "P0: BGR in + binary out"),
        "gray_erode_fast_kernel<3>": (3 * px + px, "P0: BGR in + binary out"),
        "pack_bits_kernel": (px + px // 8, "binary in + bit plane out"),
        "warp_flat_kernel<3, 0>": (3 * (px + cw * chh), "W1 NN CONTAIN: BGR in + deskewed BGR out"),
        "warp_flat_kernel<3, 1>": (3 * (px + px), "input synthesis: LINEAR, DEFAULT clip"),
        "warp_flat_kernel<1, 1>": (px + px, "input synthesis (gray)"),
        "canny_nms_kernel<3>": (3 * px + px, "H1: BGR in + state map out"),
        "canny_nms_kernel<1>": (px + px, "H1: gray in + state map out"),
        "fft_rows_kernel": (px + 16 * px, "F1 rows: u8 in + complex128 out"),
        "fft_cols_kernel": (16 * px + 4 * px, "F1 columns: complex128 in + f32 magnitude out"),
        "fft_log_kernel": (8 * px, "F1 log: f32 in + f32 out"),
        "fft_log8_kernel": (5 * px, "F1 u8: f32 in + u8 out"),
        "sweepw_kernel": (300 * px, "P2-P4: 1 B per (angle, pixel) sample, 300 angles"),
        "sweep_finalize_kernel": (300 * 4 * (W + H), "P4: counts in"),
    }


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    for i, r in enumerate(rows):
        if r and r[0] == "ID":
            hdr, start = r, i + 1
            break
    iN, iM, iV, iU = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    scale = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    per = collections.defaultdict(lambda: collections.defaultdict(list))
    for r in rows[start:]:
        if len(r) <= iV:
            continue
        name = r[iN].split("(")[0].replace("void ", "").replace("oics::", "").strip()
        per[name][r[iM]].append(float(r[iV].replace(",", "")) * scale.get(r[iU], 1.0))
    pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
    peak = float(json.load(open(pk))["hbm_gbs"]) if os.path.exists(pk) else 6650.0
    alg = algorithmic_bytes()
    out = []
    for name, m in per.items():
        t = m.get("gpu__time_duration.sum", [])
        if not t:
            continue
        us = sorted(t)[len(t) // 2]                       # median launch
        dram = None
        if m.get("dram__bytes_read.sum"):
            rd, wr = m["dram__bytes_read.sum"], m.get("dram__bytes_write.sum", [0.0] * len(t))
            dram = sorted(a + b for a, b in zip(rd, wr))[len(rd) // 2]
        row = {"kernel": name, "launches": len(t), "median_us": round(us, 2), "dram_bytes": dram}
        if name in alg:
            b, what = alg[name]
            gbs = b / (us * 1e-6) / 1e9
            row.update({"algorithmic_bytes": b, "what": what, "achieved_gbs": round(gbs, 1), "peak_gbs": peak,
                        "frac_of_hbm_peak": round(gbs / peak, 3)})
        if name == "hough_accumulate_kernel":
            cfg = os.path.join(os.path.dirname(os.path.abspath(sys.argv[1])), "configs.json")
            if os.path.exists(cfg):
                votes = json.load(open(cfg)).get("hough_accumulate", {}).get("votes")
                if votes:
                    row.update({"votes": votes, "gvotes_per_s": round(votes / (us * 1e-6) / 1e9, 3)})
        out.append(row)
    out.sort(key=lambda r: -r["median_us"] * r["launches"])
    txt = json.dumps({"card": f"{W}x{H} A4, skew {SKEW}", "note": "ncu per-launch times: cold cache, serialised", "kernels": out}, indent=1)
    if len(sys.argv) > 2:
        open(sys.argv[2], "w").write(txt)
    for r in out:
        print(f"{r['kernel']:34s} n={r['launches']:3d} {r['median_us']:10.1f} us"
              + (f"  {r['achieved_gbs']:8.1f} GB/s  {100 * r['frac_of_hbm_peak']:5.1f}% of HBM peak" if "achieved_gbs" in r else "")
              + (f"  {r['gvotes_per_s']:6.3f} G votes/s" if "gvotes_per_s" in r else ""))


if __name__ == "__main__":
    main()
