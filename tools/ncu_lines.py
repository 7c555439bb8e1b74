#!/usr/bin/env python3
"""Per source line instruction and stall-sample totals of each kernel in an .ncu-rep
(needs -lineinfo and --import-source on).  usage: ncu_lines.py rep [kernel-substring] [top]"""
import collections
import csv
import io
import subprocess
import sys


def main(rep, want="", top=40):
    out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'cuda,sass'],
                         capture_output=True, text=True).stdout
    path, fn, hdr = None, None, None
    acc = collections.defaultdict(lambda: collections.defaultdict(lambda: [0, 0, ""]))
    for r in csv.reader(io.StringIO(out)):
        if not r:
            continue
        if r[0] == 'File Path':
            path = r[1].split('/')[-1]
        elif r[0] == 'Function Name':
            fn = r[1]
        elif r[0] == 'Line No':
            hdr = r
            ie, isamp = hdr.index('Instructions Executed'), hdr.index('# Samples')
        elif hdr and r[0].strip().isdigit():
            try:
                a = acc[fn][(path, int(r[0]))]
                a[0] += int(r[ie]); a[1] += int(r[isamp]); a[2] = r[1].strip()[:90]
            except ValueError:
                pass
    for fn, lines in acc.items():
        if want not in fn:
            continue
        ti = sum(v[0] for v in lines.values()) or 1
        ts = sum(v[1] for v in lines.values()) or 1
        print('=' * 120)
        print(fn[:118], ' inst', ti, ' samples', ts)
        for (p, ln), v in sorted(lines.items(), key=lambda kv: -kv[1][1])[:top]:
            print(f'  {p:18s}:{ln:4d} inst {100 * v[0] / ti:5.1f}%  samples {100 * v[1] / ts:5.1f}%   {v[2]}')


if __name__ == '__main__':
    main(sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else "", int(sys.argv[3]) if len(sys.argv) > 3 else 40)
