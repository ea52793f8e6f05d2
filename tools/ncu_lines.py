#!/usr/bin/env python
"""Per-source-line totals (instructions executed, stall samples) of an .ncu-rep captured with
-lineinfo and --import-source on.  Usage: ncu_lines.py rep [top]"""
import csv, subprocess, sys
rep, top = sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 60
txt = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'cuda,sass'],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(txt.splitlines()))
fname, hdr, out = None, None, []
for r in rows:
    if not r:
        continue
    if r[0] == 'File Path':
        fname = r[1].split('/')[-1]
    elif r[0] == 'Line No':
        hdr = r
        ci, cs = hdr.index('Instructions Executed'), hdr.index('# Samples')
        ct = hdr.index('Thread Instructions Executed')
    elif hdr and r[0].isdigit() and len(r) > ci and r[2] == '-':      # a source line (its SASS rows carry an address)
        ins = int(r[ci]) if r[ci].isdigit() else 0
        smp = int(r[cs]) if r[cs].isdigit() else 0
        thr = int(r[ct]) if r[ct].isdigit() else 0
        if ins or smp:
            out.append((ins, smp, thr, fname, int(r[0]), r[1].strip()))
ti, ts = sum(o[0] for o in out) or 1, sum(o[1] for o in out) or 1
print('total warp-instr %d, samples %d' % (ti, ts))
for ins, smp, thr, f, ln, src in sorted(out, reverse=True)[:top]:
    print('%5.1f%% instr %5.1f%% samp  thr/instr %4.1f  %s:%d  %s' % (100.0 * ins / ti, 100.0 * smp / ts, thr / max(ins, 1), f, ln, src[:110]))
