#!/usr/bin/env python
"""Per-source-line stall samples by reason (long_sb, short_sb, wait, ...) of an .ncu-rep captured
with -lineinfo and --import-source on.  Usage: ncu_stalls.py rep [reason] [top]"""
import csv, subprocess, sys
rep = sys.argv[1]
reason = sys.argv[2] if len(sys.argv) > 2 else 'stall_long_sb'
top = int(sys.argv[3]) if len(sys.argv) > 3 else 30
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
        cr, cs = hdr.index(reason), hdr.index('# Samples')
    elif hdr and r[0].isdigit() and len(r) > cr and r[2] == '-':
        v = int(r[cr]) if r[cr].isdigit() else 0
        s = int(r[cs]) if r[cs].isdigit() else 0
        if v:
            out.append((v, s, fname, int(r[0]), r[1].strip()))
tot = sum(o[0] for o in out) or 1
print('%s: %d samples' % (reason, tot))
for v, s, f, ln, src in sorted(out, reverse=True)[:top]:
    print('%5.1f%%  (%5d of the line\'s %5d samples)  %s:%d  %s' % (100.0 * v / tot, v, s, f, ln, src[:100]))
