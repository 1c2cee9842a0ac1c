#!/usr/bin/env python
"""Summarise an .ncu-rep (raw page) and its SASS page: per-kernel counters, instruction mix
by CTA phase (split at BAR.SYNC).  Usage: tools/ncu_summary.py gpurun_out/prof.ncu-rep [out.json]"""
import collections, csv, io, json, subprocess, sys

KEYS = ['Kernel Name', 'Grid Size', 'Block Size', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'lts__t_bytes.sum',
        'lts__throughput.avg.pct_of_peak_sustained_elapsed', 'l1tex__throughput.avg.pct_of_peak_sustained_elapsed',
        'launch__registers_per_thread', 'launch__occupancy_limit_shared_mem', 'launch__occupancy_limit_registers',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
        'sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed', 'smsp__inst_executed.sum',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'sm__cycles_elapsed.max']


def run(args):
    return subprocess.run(['ncu', '-i'] + args, capture_output=True, text=True).stdout


def main():
    rep = sys.argv[1]
    rows = list(csv.reader(io.StringIO(run([rep, '--page', 'raw', '--csv']))))
    hdr, units = rows[0], rows[1]
    out = []
    for r in rows[2:]:
        d = dict(zip(hdr, r))
        out.append({k: (d[k] + ' ' + units[hdr.index(k)]).strip() for k in KEYS if k in d})
    src = list(csv.reader(io.StringIO(run([rep, '--page', 'source', '--csv']))))
    ks, cur = [], None
    for r in src:
        if r and r[0] == 'Kernel Name':
            cur = {'name': r[1], 'rows': []}
            ks.append(cur)
        elif r and r[0] == 'Address':
            cur['hdr'] = r
        elif cur is not None and r:
            cur['rows'].append(r)
    seen = set()
    for k in ks:
        if k['name'] in seen:
            continue
        seen.add(k['name'])
        h = k['hdr']
        iS, iI, iSm = h.index('Source'), h.index('Instructions Executed'), h.index('# Samples')
        tot = sum(int(r[iI]) for r in k['rows']) or 1
        totS = sum(int(r[iSm]) for r in k['rows']) or 1
        seg, segs, ops = 0, collections.defaultdict(lambda: [0, 0]), collections.Counter()
        for r in k['rows']:
            t = r[iS].split()
            op = t[1] if t[0].startswith('@') else t[0]
            if 'BAR' in r[iS]:
                seg += 1
            segs[seg][0] += int(r[iI])
            segs[seg][1] += int(r[iSm])
            ops[op.split('.')[0]] += int(r[iI])
        for o in out:
            if o['Kernel Name'].replace('(int)', '') == k['name'].replace('oics::', '').replace('(int)', ''):
                o['phase_instr_pct'] = {str(s): round(100 * v[0] / tot, 1) for s, v in segs.items()}
                o['phase_sample_pct'] = {str(s): round(100 * v[1] / totS, 1) for s, v in segs.items()}
                o['opcode_pct'] = {o_: round(100 * c / tot, 1) for o_, c in ops.most_common(12)}
    txt = json.dumps(out, indent=1)
    if len(sys.argv) > 2:
        open(sys.argv[2], 'w').write(txt)
    print(txt)


if __name__ == '__main__':
    main()
