#!/usr/bin/env python3
"""Summarise `ncu --page source --csv` output: instructions executed and stall samples per SASS address range.
usage: ncu_src_summary.py src.csv [bucket_size_in_instructions]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
ia, isrc, isamp, iexec = hdr.index('Address'), hdr.index('Source'), hdr.index('# Samples'), hdr.index('Instructions Executed')
data = [(r[isrc].strip(), int(r[isamp] or 0), int(r[iexec] or 0)) for r in rows[2:] if len(r) > iexec]
tot_s = sum(d[1] for d in data)
tot_e = sum(d[2] for d in data)
print('total samples', tot_s, 'total warp-instr', tot_e)
B = int(sys.argv[2]) if len(sys.argv) > 2 else 50
for b in range(0, len(data), B):
    ch = data[b:b + B]
    s = sum(d[1] for d in ch)
    e = sum(d[2] for d in ch)
    ops = {}
    for d in ch:
        op = d[0].split()[0] if not d[0].startswith('@') else d[0].split()[1]
        ops[op.split('.')[0]] = ops.get(op.split('.')[0], 0) + d[2]
    top = sorted(ops.items(), key=lambda kv: -kv[1])[:5]
    print(f'{b:5d}-{b + len(ch):5d} samples {100 * s / tot_s:5.1f}%  instr {100 * e / tot_e:5.1f}%  ' + ' '.join(f'{k}:{v * 100 // max(e, 1)}' for k, v in top))
