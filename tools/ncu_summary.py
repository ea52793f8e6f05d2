#!/usr/bin/env python
"""Summarise an .ncu-rep (raw page + SASS hot spots) into text.  Usage: tools_ncu_summary.py rep [kernel-index]"""
import collections, csv, re, subprocess, sys
rep = sys.argv[1]
raw = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]
keep = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread', 'launch__occupancy_limit_registers',
        'launch__occupancy_limit_shared_mem', 'launch__occupancy_limit_warps', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'smsp__thread_inst_executed_per_inst_executed.ratio', 'sm__cycles_elapsed.max', 'l1tex__t_sector_hit_rate.pct',
        'lts__t_sector_hit_rate.pct', 'smsp__issue_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum',
        'launch__grid_size', 'launch__block_size', 'l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum',
        'l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum', 'lts__t_sectors_srcunit_tex_op_read.sum',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'smsp__average_warp_latency_per_inst_issued.ratio',
        'smsp__warps_eligible.avg.per_cycle_active', 'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
        'l1tex__throughput.avg.pct_of_peak_sustained_elapsed', 'lts__throughput.avg.pct_of_peak_sustained_elapsed']
for r in rows[2:3]:
    for w in keep:
        for i, h in enumerate(hdr):
            if h == w:
                print('%-72s %s %s' % (w, r[i], units[i]))
    stall = [(float(r[i].replace(',', '')), h) for i, h in enumerate(hdr)
             if 'average_warps_issue_stalled' in h and h.endswith('_per_issue_active.ratio') and r[i]]
    print('top stalls (warps stalled per issue-active cycle):')
    for v, h in sorted(stall, reverse=True)[:8]:
        print('   %6.2f  %s' % (v, h.replace('smsp__average_warps_issue_stalled_', '').replace('_per_issue_active.ratio', '')))
sass = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'sass'], capture_output=True, text=True).stdout
rows = list(csv.reader(sass.splitlines()))
hi = [i for i, r in enumerate(rows) if r and r[0] == 'Address']
h = rows[hi[0]]
ci, cs, csm = h.index('Instructions Executed'), h.index('Source'), h.index('# Samples')
body = rows[hi[0] + 1:(hi[1] - 1 if len(hi) > 1 else len(rows))]
tot = sum(int(r[ci]) for r in body if len(r) > ci and r[ci].isdigit())
print('SASS: %d instructions, %d warp-instr executed' % (len(body), tot))
lev = collections.Counter()
samp = collections.Counter()
for r in body:
    if len(r) > ci and r[ci].isdigit():
        lev[int(r[ci])] += 1
        samp[int(r[ci])] += int(r[csm]) if r[csm].isdigit() else 0
tsamp = sum(samp.values()) or 1
for c, k in sorted(lev.items(), key=lambda kv: -kv[0] * kv[1])[:12]:
    ops = collections.Counter()
    for r in body:
        if len(r) > ci and r[ci].isdigit() and int(r[ci]) == c:
            s = re.sub(r'^@!?U?P\d+\s+', '', r[cs].strip())
            ops[s.split()[0].split('.')[0]] += 1
    print('  exec %10d x %4d instrs = %5.1f%% of instr, %5.1f%% of samples  %s' % (c, k, 100.0 * c * k / tot, 100.0 * samp[c] / tsamp, ops.most_common(8)))
