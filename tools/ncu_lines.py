#!/usr/bin/env python
"""Per-source-line totals (warp instructions executed, stall samples) of one kernel in an .ncu-rep,
by joining ncu's SASS page with `nvdisasm -g` line info of the cubin (both read here, no GPU).
usage: python tools/ncu_lines.py prof.ncu-rep libfcm.so fcm_chain chain_step_kernel [ntop]"""
import csv
import io
import os
import re
import subprocess
import sys
import tempfile


def main():
    rep, so, unit, kern = sys.argv[1:5]
    ntop = int(sys.argv[5]) if len(sys.argv) > 5 else 40
    d = tempfile.mkdtemp(dir=os.path.join(os.path.dirname(os.path.abspath(__file__)), '..', 'gpurun_out'))
    subprocess.run(['cuobjdump', '-xelf', 'all', os.path.abspath(so)], cwd=d, capture_output=True)
    cub = [f for f in os.listdir(d) if f.startswith(unit + '.')][0]
    sass = subprocess.run(['nvdisasm', '-g', '-c', os.path.join(d, cub)], capture_output=True, text=True).stdout
    line_of = {}
    cur = None
    infn = False
    for ln in sass.splitlines():
        if ln.startswith('.text.'):
            infn = kern in ln
            continue
        if not infn:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)(.*)', ln)
        if m:
            cur = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        m = re.match(r'\s*/\*([0-9a-f]{4,})\*/\s+(.*?);', ln)
        if m:
            line_of[int(m.group(1), 16)] = cur
    src = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'sass'], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(src)))
    hi = next(i for i, r in enumerate(rows) if r and r[0] == 'Address')
    hdr = rows[hi]
    ia, ie, iss = hdr.index('Address'), hdr.index('Instructions Executed'), hdr.index('Warp Stall Sampling (All Samples)')
    body = [r for r in rows[hi + 1:] if len(r) == len(hdr)]
    base = int(body[0][ia], 16)
    agg = {}
    for r in body:
        off = int(r[ia], 16) - base
        key = line_of.get(off, ('?', 0))
        a = agg.setdefault(key, [0, 0, 0])
        a[0] += int(r[ie] or 0)
        a[1] += int(r[iss] or 0)
        a[2] += 1
    ti = sum(a[0] for a in agg.values())
    ts = sum(a[1] for a in agg.values())
    srcs = {}
    print('total warp instructions %d, stall samples %d, sass instructions %d' % (ti, ts, len(body)))
    print('%-22s %8s %7s %7s %6s  %s' % ('file:line', 'inst', 'inst%', 'stall%', 'sass', 'source'))
    for key, a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:ntop]:
        f, l = key
        if f not in srcs:
            p = os.path.join(os.path.dirname(os.path.abspath(so)), 'csrc', f)
            srcs[f] = open(p).read().splitlines() if os.path.exists(p) else []
        text = srcs[f][l - 1].strip()[:90] if 0 < l <= len(srcs[f]) else ''
        print('%-22s %8d %6.1f%% %6.1f%% %6d  %s' % ('%s:%d' % (f, l), a[0], 100.0 * a[0] / max(ti, 1), 100.0 * a[1] / max(ts, 1), a[2], text))


if __name__ == '__main__':
    main()
