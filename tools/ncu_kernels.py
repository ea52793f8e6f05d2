#!/usr/bin/env python
"""One block of headline counters per profiled kernel of an .ncu-rep (raw page).  Usage: ncu_kernels.py rep"""
import csv, subprocess, sys
raw = subprocess.run(['ncu', '-i', sys.argv[1], '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
h, units = rows[0], rows[1]
keep = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum', 'l1tex__t_sector_hit_rate.pct',
        'lts__t_sector_hit_rate.pct', 'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'lts__throughput.avg.pct_of_peak_sustained_elapsed']
for r in rows[2:]:
    print('== %s   grid %s' % (r[h.index('Kernel Name')], r[h.index('Grid Size')]))
    for k in keep:
        if k in h:
            print('   %-70s %s %s' % (k, r[h.index(k)], units[h.index(k)]))
    st = sorted(((float(r[i].replace(',', '')), n) for i, n in enumerate(h)
                 if 'average_warps_issue_stalled' in n and n.endswith('_per_issue_active.ratio') and r[i]), reverse=True)[:5]
    print('   top stalls: ' + ', '.join('%s %.2f' % (n.replace('smsp__average_warps_issue_stalled_', '').replace('_per_issue_active.ratio', ''), v) for v, n in st))
