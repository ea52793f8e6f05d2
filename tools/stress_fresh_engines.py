import sys, os; sys.path.insert(0, '/root/repo')
import numpy as np
from tests import oracle_lib as O
from shellac_b200 import _native as N
from shellac_b200.engine import Engine
g = O.load_golden()
msgs = [w['msg'] for w in g['whole']]
o = None
for it in range(150):
    eng = Engine(device=0, max_msgs=1<<20, max_bytes=(1<<30), max_records=1<<25)
    ar = eng.pack(msgs)
    n = ar.n
    if o is None:
        o = O.orc_parse(ar)
    po = eng.parse(ar)
    base = po.hdr_base[:n].astype(np.int64); cnt = po.n_headers[:n].astype(np.int64)
    order = np.argsort(base, kind='stable')
    b, c = base[order], cnt[order]
    nz = c > 0
    b, c, idxs = b[nz], c[nz], order[nz]
    overlap = np.flatnonzero(b[1:] < b[:-1] + c[:-1])
    bad = []
    for i in range(n):
        x = po.hdr[int(base[i]):int(base[i]) + int(cnt[i])]
        y = o.hdr[int(o.hdr_base[i]):int(o.hdr_base[i]) + int(o.n_headers[i])]
        if len(x) != len(y) or (x != y).any():
            bad.append(i)
    if bad or len(overlap):
        print('iter', it, 'overlaps', len(overlap), 'bad msgs', bad, 'max end', int((b + c).max()), 'used', int(po.hdr_used[0]))
        for i in bad[:3]:
            print('  msg', i, 'blk', i // 128, 'code', N.CODE_NAMES.get(int(po.status[i]) & 0xff), 'base', int(base[i]), 'n', int(cnt[i]),
                  'gpu', po.hdr[int(base[i]):int(base[i]) + int(cnt[i])].tolist(), 'orc', o.hdr[int(o.hdr_base[i]):int(o.hdr_base[i]) + int(o.n_headers[i])].tolist(), repr(msgs[i][:70]))
        for k in overlap[:3]:
            print('  overlap between msg', int(idxs[k]), '(base %d n %d)' % (b[k], c[k]), 'and msg', int(idxs[k + 1]), '(base %d n %d)' % (b[k + 1], c[k + 1]))
        break
    eng.close()
print('done')
