#!/usr/bin/env python
"""Instruction-cache footprint of one kernel from an .ncu-rep captured with --import-source on (read here, no GPU).

For every SASS instruction the source page gives how often it was executed; this prints
  * how many 128-byte instruction lines (the L0 line; the L1.5 instruction cache holds 32 KB) cover 50..99 % of the
    executed warp instructions,
  * how many lines hold an instruction that runs at least x times per proposal ("the walk" of one proposal),
  * the same per source function of csrc/fcm_chain.cu (innermost inlined function of each SASS instruction),
  * 4 KB address bins with their share of executed instructions and of `no_instruction` stall samples.
usage: python tools/ncu_footprint.py report.ncu-rep <proposals in the launch> [source.cu]"""
import collections
import csv
import io
import re
import subprocess
import sys


def functions(path):
    """[(first line, name)] of the top-level function definitions of a .cu file (brace depth 0)"""
    out, depth, sig, sig_line = [], 0, "", 0
    for ln, t in enumerate(open(path), 1):
        code = t.split("//")[0]
        if depth == 0:
            s = code.strip()
            if s and not s.startswith("#"):
                if not sig:
                    sig_line = ln
                sig += " " + s
                if "{" in s or s.endswith(";"):
                    if "(" in sig and "{" in s:
                        head = sig.split("(")[0]
                        m = re.findall(r"[A-Za-z_]\w*", head)
                        if m and m[-1] not in ("if", "for", "while", "switch"):
                            out.append((sig_line, m[-1]))
                    sig = ""
        depth += code.count("{") - code.count("}")
    return out


def main():
    rep, proposals = sys.argv[1], float(sys.argv[2])
    src_path = sys.argv[3] if len(sys.argv) > 3 else "flag_complex_mcmc_b200/csrc/fcm_chain.cu"
    fns = functions(src_path)
    txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                         capture_output=True, text=True).stdout
    rows = [[("" if c == "-" else c) for c in r] for r in csv.reader(io.StringIO(txt))]
    hi = next(i for i, r in enumerate(rows) if r and r[0] == "Line No")
    hdr = rows[hi]
    i_ex = hdr.index("Instructions Executed")
    i_ni = hdr.index("stall_no_inst") if "stall_no_inst" in hdr else None
    ins, cur, fileidx = {}, None, -1
    for r in rows[hi:]:
        if r and r[0] == "Line No":
            fileidx += 1
            continue
        if len(r) != len(hdr):
            continue
        if r[0]:
            cur = (fileidx, int(r[0]))
        if not r[2]:
            continue
        try:
            a = int(r[2], 16)
        except ValueError:
            continue
        ins[a] = (int(r[i_ex] or 0), int(r[i_ni] or 0) if i_ni is not None else 0, cur)

    def func(c):
        f, ln = c
        if f != 0:
            return "(cuda headers: shuffles, votes, atomics)"
        name = "?"
        for s, n in fns:
            if s <= ln:
                name = n
            else:
                break
        return name

    tot = sum(v[0] for v in ins.values())
    tni = sum(v[1] for v in ins.values())
    print("# %s: %d SASS instructions (%.0f KB), %d executed warp instructions, %.0f per proposal" % (
        rep.split("/")[-1], len(ins), len(ins) * 16 / 1024, tot, tot / proposals))
    lines = collections.defaultdict(list)
    for a, v in ins.items():
        lines[a >> 7].append(v[0])
    lw = sorted((sum(v) for v in lines.values()), reverse=True)
    acc, marks, mi = 0, [0.5, 0.8, 0.9, 0.95, 0.99], 0
    for k, s in enumerate(lw):
        acc += s
        while mi < len(marks) and acc >= marks[mi] * tot:
            print("%2.0f %% of the executed instructions lie in %4d lines of 128 B = %5.1f KB" % (marks[mi] * 100, k + 1, (k + 1) / 8))
            mi += 1
    for th in (1.0, 0.5, 0.2, 0.05):
        nl = sum(1 for v in lines.values() if max(v) >= th * proposals)
        print("lines with an instruction executed >= %.2f x per proposal: %4d = %5.1f KB" % (th, nl, nl / 8))
    per = collections.defaultdict(lambda: [set(), 0, 0])
    for a, (ex, ni, c) in ins.items():
        d = per[func(c)]
        d[1] += ex
        d[2] += ni
        if ex >= 0.2 * proposals:
            d[0].add(a >> 7)
    print("-- per source function: lines walked (>= 0.2 x per proposal; a line can be shared), share of executed instructions, share of no_instruction samples")
    for f, d in sorted(per.items(), key=lambda kv: -len(kv[1][0]))[:28]:
        if len(d[0]) or d[1] > 0.005 * tot:
            print("%-44s %4d lines %5.1f KB   %5.1f %% inst   %5.1f %% no_inst" % (f, len(d[0]), len(d[0]) / 8, 100.0 * d[1] / tot, 100.0 * d[2] / max(tni, 1)))
    print("-- 4 KB address bins (share of executed instructions / of no_instruction samples >= 0.5 %)")
    addrs = sorted(ins)
    for b in range(0, len(addrs), 256):
        seg = addrs[b:b + 256]
        d = sum(ins[a][0] for a in seg)
        n = sum(ins[a][1] for a in seg)
        if d > 0.005 * tot or n > 0.005 * tni:
            top = collections.Counter()
            for a in seg:
                top[func(ins[a][2])] += ins[a][0]
            print("%6d  %5.1f %%  %5.1f %%   %s" % (b, 100.0 * d / tot, 100.0 * n / max(tni, 1), ", ".join(k for k, _ in top.most_common(3))))


if __name__ == "__main__":
    main()
