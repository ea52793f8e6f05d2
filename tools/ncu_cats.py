#!/usr/bin/env python
"""Warp-instructions and stall samples of an .ncu-rep per FUNCTION / kernel stage of
parse_kernel.cuh (each SASS instruction counted once, under the innermost frame that lies in
that file).  Stages inside the kernel are the '// ---- ' comments.  Usage: ncu_cats.py rep [src]"""
import collections, csv, re, subprocess, sys
rep = sys.argv[1]
txt = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'cuda,sass'], capture_output=True, text=True).stdout
rows = list(csv.reader(txt.splitlines()))
# line -> stage map from the source file on disk (profile right after building, so that it matches)
src_path = sys.argv[2] if len(sys.argv) > 2 else 'shellac_b200/csrc/parse_kernel.cuh'
marks = []
for ln, line in enumerate(open(src_path), 1):
    src = line.strip()
    m = re.match(r'(?:__host__ )?__(?:device|global)__ (?:__\w+__ )*(?:inline )?(?:[\w:<>]+[ *&]+)+?(\w+)\s*\(', src)
    if m and m.group(1) != '__launch_bounds__':
        marks.append((ln, m.group(1)))
    elif '__global__' in src:
        marks.append((ln, 'k_route: head'))
    elif src.startswith('// ---- '):
        marks.append((ln, 'k_route: ' + src[8:60]))
marks = sorted(set(marks))
def stage(line):
    name = 'parse_kernel.cuh (top)'
    for ln, n in marks:
        if ln <= line:
            name = n
    return name
agg, samp, seen = collections.Counter(), collections.Counter(), set()
fname = None
for r in rows:
    if not r:
        continue
    if r[0] == 'File Path':
        fname = r[1].split('/')[-1]
    elif r[0] == 'Line No':
        ci, cs = r.index('Instructions Executed'), r.index('# Samples')
    elif r[0].isdigit() and len(r) > 2 and r[2] == '-':
        cur = (fname, int(r[0]))
    elif r[0] == '' and len(r) > ci and r[2].startswith('0x') and r[ci].isdigit():
        if r[2] in seen:
            continue
        seen.add(r[2])
        key = stage(cur[1]) if cur[0] == 'parse_kernel.cuh' else cur[0]
        agg[key] += int(r[ci])
        samp[key] += int(r[cs]) if r[cs].isdigit() else 0
tot, ts = sum(agg.values()) or 1, sum(samp.values()) or 1
print('total warp-instr %d' % tot)
for k, v in agg.most_common():
    if v * 1000 > tot:
        print('%6.2f%% instr %6.2f%% samples  %s' % (100.0 * v / tot, 100.0 * samp[k] / ts, k))
