#!/usr/bin/env python
"""Aggregate an ncu source-page CSV (--print-source cuda,sass) per CUDA source line:
instructions executed, stall samples and the dominant stall reason.  usage: ncu_lines2.py file.csv [n]"""
import csv, sys, collections
rows = [[("" if c == "-" else c) for c in r] for r in csv.reader(open(sys.argv[1]))]
nshow = int(sys.argv[2]) if len(sys.argv) > 2 else 60
hi = next(i for i, r in enumerate(rows) if r and r[0] == 'Line No')
hdr = rows[hi]
iS = hdr.index('Warp Stall Sampling (All Samples)'); iI = hdr.index('Instructions Executed'); iL = 0; iSrc = 1; iAddr = 2; iSass = 3
stall_cols = [i for i, h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
per = collections.OrderedDict()
cur = None
for r in rows[hi + 1:]:
    if len(r) != len(hdr):
        continue
    if r[iL] == 'Line No':
        continue
    if r[iL]:
        cur = (int(r[iL]), r[iSrc].strip())
    if cur is None or not r[iAddr]:
        continue
    d = per.setdefault(cur, {'inst': 0, 'smp': 0, 'n': 0, 'st': collections.Counter(), 'local': 0})
    d['inst'] += int(r[iI] or 0); d['smp'] += int(r[iS] or 0); d['n'] += 1
    if 'LDL' in r[iSass] or 'STL' in r[iSass]:
        d['local'] += int(r[iI] or 0)
    for i in stall_cols:
        d['st'][hdr[i]] += int(r[i] or 0)
ti = sum(d['inst'] for d in per.values()); ts = sum(d['smp'] for d in per.values())
print('total warp inst', ti, 'samples', ts, 'local inst', sum(d['local'] for d in per.values()))
print('-- by instructions')
for (ln, src), d in sorted(per.items(), key=lambda kv: -kv[1]['inst'])[:nshow]:
    top = d['st'].most_common(1)[0][0] if d['st'] else ''
    print('%5d %5.1f%% inst %5.1f%% smp %4d sass local=%-9d %-14s %s' % (ln, 100.0 * d['inst'] / ti, 100.0 * d['smp'] / max(ts, 1), d['n'], d['local'], top, src[:100]))
print('-- by stall samples')
for (ln, src), d in sorted(per.items(), key=lambda kv: -kv[1]['smp'])[:nshow]:
    top = d['st'].most_common(2)
    print('%5d %5.1f%% smp %5.1f%% inst %-30s %s' % (ln, 100.0 * d['smp'] / max(ts, 1), 100.0 * d['inst'] / ti, ','.join('%s=%d' % kv for kv in top), src[:100]))
