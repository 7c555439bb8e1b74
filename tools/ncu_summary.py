#!/usr/bin/env python3
"""Summarise an .ncu-rep: key metrics per kernel, opcode mix and stall reasons (needs ncu on PATH)."""
import collections
import csv
import io
import re
import subprocess
import sys

KEYS = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'lts__t_bytes.sum',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread', 'launch__occupancy_limit_registers',
        'launch__occupancy_limit_shared_mem', 'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
        'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active', 'sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'l1tex__data_pipe_lsu_wavefronts.sum', 'l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed',
        'l1tex__lsu_writeback_active.avg.pct_of_peak_sustained_active', 'l1tex__t_sector_hit_rate.pct',
        'lts__t_sector_hit_rate.pct', 'sm__cycles_elapsed.avg', 'sm__cycles_active.avg', 'smsp__cycles_active.avg']


def run(args):
    return subprocess.run(['ncu'] + args, capture_output=True, text=True).stdout


def main(rep):
    raw = list(csv.reader(io.StringIO(run(['-i', rep, '--page', 'raw', '--csv']))))
    hdr, units = raw[0], raw[1]
    for r in raw[2:]:
        print('=' * 100)
        print(r[hdr.index('Kernel Name')][:110])
        for k in KEYS:
            if k in hdr:
                i = hdr.index(k)
                print(f'  {k:75s} {r[i]:>18s} {units[i]}')
    src = list(csv.reader(io.StringIO(run(['-i', rep, '--page', 'source', '--csv']))))
    ks, cur = [], None
    for r in src:
        if r and r[0] == 'Kernel Name':
            cur = {'name': r[1], 'rows': []}
            ks.append(cur)
        elif r and r[0] == 'Address':
            cur['hdr'] = r
        elif cur is not None and r:
            cur['rows'].append(r)
    for k in ks:
        h = k['hdr']
        ie = h.index('Instructions Executed')
        stallcols = [i for i, x in enumerate(h) if x.startswith('stall_') and 'Not Issued' not in x]
        ops, stalls, tot = collections.Counter(), collections.Counter(), 0
        for r in k['rows']:
            try:
                n = int(r[ie])
            except ValueError:
                continue
            m = re.match(r'\s*(@!?U?P\d+\s+)?([A-Z0-9_.]+)', r[1])
            ops[m.group(2).split('.')[0] if m else '?'] += n
            tot += n
            for i in stallcols:
                try:
                    stalls[h[i]] += int(r[i])
                except ValueError:
                    pass
        print('-' * 100)
        print(k['name'][:110], ' warp-instructions', tot)
        print('  opcodes: ' + '  '.join(f'{o} {100 * n / tot:.1f}%' for o, n in ops.most_common(22)))
        ts = sum(stalls.values()) or 1
        print('  stalls : ' + '  '.join(f'{s[6:]} {100 * n / ts:.1f}%' for s, n in stalls.most_common(9)))


if __name__ == '__main__':
    main(sys.argv[1])
