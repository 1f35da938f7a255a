#!/usr/bin/env python
"""Summarise an .ncu-rep (read here, no GPU): key raw metrics + top source lines by stall samples.
usage: python tools/ncu_summary.py gpurun_out/prof.ncu-rep [nlines] [key=value ...]
key=value pairs are written as a '# key=value ...' header line (bench.py reads e.g. chains=4096 proposals=500 from it)."""
import csv
import io
import subprocess
import sys

WANT = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'launch__registers_per_thread', 'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'smsp__inst_executed.sum',
        'sm__inst_executed_pipe_lsu.sum.pct_of_peak_sustained_active', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'l1tex__t_sector_hit_rate.pct', 'lts__t_sector_hit_rate.pct', 'smsp__thread_inst_executed_per_inst_executed.ratio',
        'launch__waves_per_multiprocessor', 'sass__inst_executed_local_loads', 'sass__inst_executed_local_stores',
        'launch__grid_size', 'launch__block_size', 'launch__occupancy_limit_shared_mem', 'launch__occupancy_limit_registers',
        'sm__cycles_elapsed.avg', 'launch__occupancy_limit_warps', 'launch__occupancy_limit_blocks',
        'sm__maximum_warps_per_active_cycle_pct', 'launch__shared_mem_per_block_dynamic', 'lts__t_bytes.sum', 'l1tex__t_bytes.sum',
        'smsp__warps_eligible.avg.per_cycle_active', 'smsp__issue_inst0.avg.pct_of_peak_sustained_active']


def main():
    rep = sys.argv[1]
    nl = int(sys.argv[2]) if len(sys.argv) > 2 and "=" not in sys.argv[2] else 25
    kv = [a for a in sys.argv[2:] if "=" in a]
    if kv:
        print("# " + " ".join(kv))
    raw = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    for vals in rows[2:]:
        print('== kernel:', vals[hdr.index('Kernel Name')][:80] if 'Kernel Name' in hdr else '')
        print('metric,unit,value')
        for h, u, v in zip(hdr, units, vals):
            if h in WANT or h.startswith('smsp__average_warps_issue_stalled') and h.endswith('_per_issue_active.ratio'):
                print('%s,%s,%s' % (h, u, v))
    src = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'sass'], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(src)))
    # find header row
    hi = next(i for i, r in enumerate(rows) if r and r[0] == 'Address')
    hdr = rows[hi]
    ci = {k: hdr.index(k) for k in ('Source', 'Warp Stall Sampling (All Samples)', 'Instructions Executed') if k in hdr}
    body = [r for r in rows[hi + 1:] if len(r) == len(hdr)]
    tot = sum(int(r[ci['Warp Stall Sampling (All Samples)']] or 0) for r in body)
    toti = sum(int(r[ci['Instructions Executed']] or 0) for r in body)
    print('== total stall samples', tot, 'total warp instructions', toti)
    stall_cols = [i for i, h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
    agg = {hdr[i]: sum(int(r[i] or 0) for r in body) for i in stall_cols}
    print('== stall reasons (all samples):', ', '.join('%s=%.1f%%' % (k, 100.0 * v / max(tot, 1)) for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]))
    body_sorted = sorted(enumerate(body), key=lambda ir: -int(ir[1][ci['Warp Stall Sampling (All Samples)']] or 0))
    print('== top SASS lines by stall samples (idx, samples%, executed, sass)')
    for idx, r in body_sorted[:nl]:
        smp = int(r[ci['Warp Stall Sampling (All Samples)']] or 0)
        top = max(stall_cols, key=lambda i: int(r[i] or 0))
        print('%5d %5.1f%% %10s  %-60s %s' % (idx, 100.0 * smp / max(tot, 1), r[ci['Instructions Executed']], r[ci['Source']].strip()[:60], hdr[top]))


if __name__ == '__main__':
    main()
