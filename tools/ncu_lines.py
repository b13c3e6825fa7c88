#!/usr/bin/env python
"""Per-source-line summary of an ncu report (needs -lineinfo + --import-source on).

usage: tools/ncu_lines.py report.ncu-rep [min_pct]
Prints, for every CUDA source line above min_pct of executed warp instructions or of
stall samples: file:line, % instructions, % samples, dominant stall reasons, source text.
"""
import csv, io, subprocess, sys

rep = sys.argv[1]
min_pct = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
txt = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--print-source', 'cuda,sass', '--csv'],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
cur_file, hdr, lines = None, None, []
for r in rows:
    if not r:
        continue
    if r[0] == 'File Path':
        cur_file = r[1].split('/')[-1]
    elif r[0] == 'Line No':
        hdr = r
    elif hdr and len(r) == len(hdr) and r[2] == '-' and r[0].isdigit():
        d = dict(zip(hdr[4:], r[4:]))
        lines.append((cur_file, int(r[0]), r[1].strip(), d))
tot_i = sum(int(d['Instructions Executed']) for *_, d in lines) or 1
tot_s = sum(int(d['# Samples']) for *_, d in lines) or 1
print(f'total warp instructions {tot_i}, samples {tot_s}')
stalls = [k for k in lines[0][3] if k.startswith('stall_') and 'Not Issued' not in k]
for f, n, src, d in lines:
    pi = 100 * int(d['Instructions Executed']) / tot_i
    ps = 100 * int(d['# Samples']) / tot_s
    if pi >= min_pct or ps >= min_pct:
        top = sorted(((int(d[k]), k[6:]) for k in stalls), reverse=True)[:2]
        ts = ','.join(f'{k}:{v}' for v, k in top if v)
        print(f'{f}:{n:<4d} inst {pi:5.1f}%  smp {ps:5.1f}%  [{ts}]  {src[:90]}')
