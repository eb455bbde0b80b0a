#!/usr/bin/env python
"""Key metrics + stall ranking of an .ncu-rep (first profiled launch).  usage: tools/ncu_summary.py file.ncu-rep [out.csv]"""
import csv, io, subprocess, sys
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, r = rows[0], rows[1], rows[2]
want = ['Kernel Name', 'Grid Size', 'Block Size', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'launch__registers_per_thread', 'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem',
        'launch__waves_per_multiprocessor', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'smsp__inst_executed.sum', 'smsp__cycles_active.avg', 'sm__cycles_elapsed.max',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active',
        'l1tex__t_sector_hit_rate.pct', 'lts__t_sector_hit_rate.pct', 'smsp__thread_inst_executed_per_inst_executed.ratio',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'sm__inst_executed_pipe_fp64.sum', 'sm__inst_executed_pipe_xu.sum']
out = []
for w in want:
    if w in hdr:
        i = hdr.index(w); out.append((w, units[i], r[i]))
stall = [(h, float(r[i].replace(',', ''))) for i, h in enumerate(hdr)
         if 'smsp__average_warp' in h and 'issue_stalled' in h and 'not_issued' not in h and r[i]]
for h, v in sorted(stall, key=lambda x: -x[1])[:12]:
    out.append((h.replace('smsp__average_warps_issue_stalled_', 'stall_').replace('_per_issue_active.ratio', ''), 'warps/issue', v))
for o in out:
    print(*o)
if len(sys.argv) > 2:
    with open(sys.argv[2], 'w') as f:
        w = csv.writer(f); w.writerow(['metric', 'unit', 'value']); w.writerows(out)
