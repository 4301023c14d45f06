#!/usr/bin/env python3
"""Prints the headline metrics and the hottest source lines of each kernel in an .ncu-rep."""
import csv, subprocess, sys
rep = sys.argv[1]
kern = sys.argv[2] if len(sys.argv) > 2 else None
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr = rows[0]
want = ['Kernel Name', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'launch__registers_per_thread', 'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'smsp__inst_executed.sum', 'smsp__thread_inst_executed_per_inst_executed.ratio',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'lts__t_sector_hit_rate.pct']
want += [h for h in hdr if h.startswith('smsp__average_warps_issue_stalled') and h.endswith('per_issue_active.ratio')]
for r in rows[2:]:
    if kern and kern not in r[hdr.index('Kernel Name')]:
        continue
    print('=' * 60)
    stalls = []
    for w in want:
        if w in hdr:
            v = r[hdr.index(w)]
            if w.startswith('smsp__average_warps_issue_stalled'):
                try:
                    stalls.append((float(v), w.replace('smsp__average_warps_issue_stalled_', '').replace('_per_issue_active.ratio', '')))
                except ValueError:
                    pass
            else:
                print(f'{w:70s} {v[:60]}')
    print('stalls:', ', '.join(f'{n}={v:.2f}' for v, n in sorted(stalls, reverse=True)[:7]))
if kern:
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kern, "--print-source", "sass,cuda"],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(src.splitlines()))
    files = {}
    cur = None
    hi = None
    agg = []
    for i, r in enumerate(rows):
        if r and r[0] == "File Path":
            cur = r[1].split('/')[-1]
        if r and r[0] == "Line No":
            hdr2 = r
            isamp = hdr2.index("# Samples"); iinst = hdr2.index("Instructions Executed"); ithr = hdr2.index("Thread Instructions Executed")
            continue
        if cur and len(r) > 10 and r[2] == '-':
            try:
                agg.append((int(r[iinst]), int(r[isamp]), int(r[ithr]), cur, r[0], r[1][:100]))
            except (ValueError, NameError):
                pass
    tot = sum(a[0] for a in agg) or 1
    tots = sum(a[1] for a in agg) or 1
    print("total warp-inst", tot, "samples", tots)
    for a in sorted(agg, reverse=True)[:int(sys.argv[3]) if len(sys.argv) > 3 else 25]:
        print(f"{a[0]/tot*100:5.1f}% inst {a[1]/tots*100:5.1f}% samp thr/inst={a[2]/max(a[0],1):5.1f} {a[3]}:{a[4]:>4s}: {a[5]}")
