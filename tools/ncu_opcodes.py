#!/usr/bin/env python
"""Executed warp instructions per SASS opcode for one kernel of an ncu report (needs
--import-source on).  usage: tools/ncu_opcodes.py report.ncu-rep [kernel-name-substring]"""
import csv, io, subprocess, sys
from collections import Counter

rep = sys.argv[1]
want = sys.argv[2] if len(sys.argv) > 2 else ''
txt = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--print-source', 'sass', '--csv'],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
hdr, kernel, done = None, None, set()
ops, total = Counter(), 0
for r in rows:
    if not r:
        continue
    if r[0] == 'Kernel Name':
        kernel = r[1]
        hdr = None
        continue
    if 'Source' in r and 'Instructions Executed' in r:
        hdr = r
        if kernel in done:
            hdr = None  # first launch of each kernel only
        done.add(kernel)
        continue
    if hdr and len(r) == len(hdr) and (want in (kernel or '')):
        d = dict(zip(hdr, r))
        try:
            n = int(d['Instructions Executed'])
        except ValueError:
            continue
        src = d['Source'].strip()
        if src.startswith('@'):
            src = src.split(None, 1)[1] if ' ' in src else src
        op = src.split()[0].split('.')[0] if src else '?'
        full = '.'.join(src.split()[0].split('.')[:2]) if src else '?'
        ops[full] += n
        total += n
print(f'kernel filter "{want}": {total} warp instructions')
for op, n in ops.most_common(45):
    print(f'  {op:20s} {n:14d}  {100 * n / total:5.1f}%')
