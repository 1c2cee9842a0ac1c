#!/usr/bin/env python
"""Per-source-line instruction and stall-sample shares of an .ncu-rep captured with --import-source on
(-lineinfo builds).  Usage: tools/ncu_lines.py rep.ncu-rep [top_n] [kernel-substring]"""
import csv, io, subprocess, sys, collections
rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
txt = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'cuda,sass'], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
lines = []   # (file, line, src, instr, samples, shared_wavefronts, excessive)
cur_file = None
hdr = None
for r in rows:
    if not r:
        continue
    if r[0] == 'File Path':
        cur_file = r[1].split('/')[-1]
    elif r[0] == 'Line No':
        hdr = r
        iI, iS = hdr.index('Instructions Executed'), hdr.index('# Samples')
        iW = hdr.index('L1 Wavefronts Shared') if 'L1 Wavefronts Shared' in hdr else iI
        iX = hdr.index('L1 Wavefronts Shared Excessive') if 'L1 Wavefronts Shared Excessive' in hdr else iI
    elif hdr and r[0].isdigit():
        n=lambda v: int(v) if v.lstrip("-").isdigit() else 0; lines.append((cur_file, int(r[0]), r[1].strip(), n(r[iI]), n(r[iS]), n(r[iW]), n(r[iX])))
totI = sum(l[3] for l in lines) or 1
totS = sum(l[4] for l in lines) or 1
print(f"total warp instructions {totI:,}  samples {totS:,}")
for f, ln, src, ins, smp, wf, ex in sorted(lines, key=lambda l: -l[3])[:top]:
    print(f"{100*ins/totI:5.1f}% ins {100*smp/totS:5.1f}% smp  wf {wf:>11,} ex {ex:>10,}  {f}:{ln}  {src[:110]}")
