#!/usr/bin/env python3
"""Top warp-stall reasons (stalled warps per issue-active cycle) of every kernel in an ncu report: ncu_stalls.py report.ncu-rep"""
import csv
import subprocess
import sys


def main(path):
    out = subprocess.run(['ncu', '-i', path, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    h = rows[0]
    kn = h.index('Kernel Name')
    cols = [j for j, x in enumerate(h) if x.startswith('smsp__average_warps_issue_stalled_') and x.endswith('_per_issue_active.ratio')]
    for r in rows[2:]:
        st = sorted(((float(r[j].replace(',', '') or 0), h[j][len('smsp__average_warps_issue_stalled_'):-len('_per_issue_active.ratio')]) for j in cols), reverse=True)
        print('%-60s %s' % (r[kn][:60], '  '.join('%s=%.2f' % (n, v) for v, n in st[:6])))


if __name__ == '__main__':
    main(sys.argv[1])
