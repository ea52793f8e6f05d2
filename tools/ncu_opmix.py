#!/usr/bin/env python
"""Dynamic opcode mix (warp-instructions executed per SASS opcode) of an .ncu-rep.  Usage: ncu_opmix.py rep"""
import collections, csv, re, subprocess, sys
txt = subprocess.run(['ncu', '-i', sys.argv[1], '--page', 'source', '--csv', '--print-source', 'sass'], capture_output=True, text=True).stdout
rows = list(csv.reader(txt.splitlines()))
hi = [i for i, r in enumerate(rows) if r and r[0] == 'Address'][0]
h = rows[hi]
ci, cs = h.index('Instructions Executed'), h.index('Source')
mix = collections.Counter()
for r in rows[hi + 1:]:
    if len(r) > ci and r[ci].isdigit():
        op = re.sub(r'^@!?U?P\d+\s+', '', r[cs].strip()).split()[0].split('.')[0]
        mix[op] += int(r[ci])
tot = sum(mix.values())
print('total', tot)
for op, c in mix.most_common(32):
    print('%-10s %6.2f%%' % (op, 100.0 * c / tot))
