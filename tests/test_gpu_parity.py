"""GPU parity tests proper (-m gpu): every call goes through the C ABI (ctypes ->
libshellac_b200.so -> sm_100a kernels) and is compared with (a) the golden vectors the unmodified
reference produced and (b) the CPU oracle on the same seeded inputs.  Bar: bit-exact."""
import hashlib

import numpy as np
import pytest

from shellac_b200 import _native as N
from shellac_b200.batch import Arena, MessageView
from tests import oracle_lib as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope='module')
def eng():
    from shellac_b200 import build as B
    from shellac_b200.engine import Engine
    B.build()
    e = Engine(device=0, max_msgs=1 << 20, max_bytes=(1 << 30) + (1 << 28), max_records=1 << 25)
    yield e
    e.close()


# ------------------------------------------------------------------------------- helpers
def _records(out, n):
    """Header records of messages [0,n) in message order, as one uint64 array (+ per-msg counts)."""
    cnt = out.n_headers[:n].astype(np.int64)
    base = out.hdr_base[:n].astype(np.int64)
    total = int(cnt.sum())
    if total == 0:
        return np.zeros(0, np.uint64)
    excl = np.cumsum(cnt) - cnt
    idx = np.repeat(base - excl, cnt) + np.arange(total)
    return out.hdr.view(np.uint64)[idx]


def _describe(ar, i, g, o):
    m = ar.message(i)
    fields = [f for f, _ in N.PARSE_FIELDS if f != 'hdr_base']
    return {'i': i, 'msg': m[:160], 'len': len(m),
            'gpu': {f: int(getattr(g, f)[i]) for f in fields}, 'orc': {f: int(getattr(o, f)[i]) for f in fields},
            'gpu_code': N.CODE_NAMES.get(int(g.status[i]) & 0xff), 'orc_code': N.CODE_NAMES.get(int(o.status[i]) & 0xff)}


def assert_parse_equal(ar, g, o, what):
    n = ar.n
    bad = np.zeros(n, bool)
    for f, _ in N.PARSE_FIELDS:
        if f == 'hdr_base':
            continue
        bad |= getattr(g, f)[:n] != getattr(o, f)[:n]
    if bad.any():
        w = np.flatnonzero(bad)
        raise AssertionError('%s: %d/%d messages differ; first: %r' % (what, len(w), n, [_describe(ar, int(i), g, o) for i in w[:4]]))
    rg, ro = _records(g, n), _records(o, n)
    if len(rg) != len(ro) or (rg != ro).any():
        cnt = o.n_headers[:n].astype(np.int64)
        owner = np.repeat(np.arange(n), cnt)
        k = int(np.flatnonzero(rg != ro)[0])
        i = int(owner[k])
        b_g, b_o, c = int(g.hdr_base[i]), int(o.hdr_base[i]), int(cnt[i])
        raise AssertionError('%s: header records differ, first at message %d: %r\n gpu=%r\n orc=%r' % (
            what, i, ar.message(i)[:300], g.hdr[b_g:b_g + c].tolist(), o.hdr[b_o:b_o + c].tolist()))
    assert int(g.hdr_used[0]) == int(o.hdr_used[0]), what


def assert_frame_equal(ar, g, o, what):
    n = ar.n
    for f, _ in N.FRAME_FIELDS:
        if f == 'chunk_base':
            continue
        a, b = getattr(g, f)[:n], getattr(o, f)[:n]
        if (a != b).any():
            i = int(np.flatnonzero(a != b)[0])
            raise AssertionError('%s: frame field %s differs at %d: gpu=%d orc=%d msg=%r' % (
                what, f, i, int(a[i]), int(b[i]), ar.message(i)[:300]))
    for i in np.flatnonzero(o.n_chunks[:n]):
        c = int(o.n_chunks[i])
        x = g.chunk[int(g.chunk_base[i]):int(g.chunk_base[i]) + c]
        y = o.chunk[int(o.chunk_base[i]):int(o.chunk_base[i]) + c]
        assert (x == y).all(), (what, int(i), x.tolist(), y.tolist())


def scatter_pack(eng, msgs, rng):
    """Pack with random 0..15-byte gaps so message starts hit every 16-byte alignment, and fill
    the gaps with bytes that would confuse a parser reading outside [off, off+len)."""
    msgs = [m.encode('latin-1') if isinstance(m, str) else m for m in msgs]
    gaps = rng.integers(0, 16, len(msgs))
    junk = b'\r\n\r\n: \r\n \r\n\r\n: '
    parts, off, pos = [], [], 0
    for m, gp in zip(msgs, gaps):
        parts.append(junk[:gp]); pos += gp
        off.append(pos); parts.append(m); pos += len(m)
    blob = b''.join(parts) + junk
    data = eng.pinned(len(blob) + 64, np.uint8)
    data[:len(blob)] = np.frombuffer(blob, np.uint8)
    a_off, a_len = eng.pinned(len(msgs), np.uint64), eng.pinned(len(msgs), np.uint32)
    a_off[:] = off
    a_len[:] = [len(m) for m in msgs]
    return Arena(data, a_off, a_len)


# ------------------------------------------------------------------------------- golden
@pytest.mark.parametrize('force_scalar', [False, True])
def test_golden_whole_messages(eng, golden, force_scalar):
    """Config 1: the reference's own corpus (test/HttpParserTests.py literals + seeded poison)."""
    eng.set_force_scalar(force_scalar)
    try:
        ar = eng.pack([w['msg'] for w in golden['whole']])
        po = eng.parse(ar)
        fo = eng.frame(ar, po, reuse_arena=True)
    finally:
        eng.set_force_scalar(False)
    bad = []
    for i, w in enumerate(golden['whole']):
        v = MessageView(ar.message(i), po, i, fo, text=True)
        d = O.diff_observation(v, w['obs'])
        if d:
            bad.append((w['id'], w['msg'][:100], d))
    assert not bad, (len(bad), bad[:5])


def test_golden_vs_oracle_soa_all_alignments(eng, golden):
    rng = np.random.default_rng(1)
    msgs = [w['msg'] for w in golden['whole']]
    for rep in range(3):
        ar = scatter_pack(eng, msgs, rng)
        g = eng.parse(ar)
        o = O.orc_parse(ar)
        assert_parse_equal(ar, g, o, 'golden scatter %d' % rep)
        assert_frame_equal(ar, eng.frame(ar, g, reuse_arena=True), O.orc_frame(ar, o), 'golden frame %d' % rep)


def test_unordered_arena_takes_the_per_lane_scan(eng, golden):
    """k_route scans a group of 32 messages as one span of bytes when they lie in index order;
    an arena indexed back to front (or with overlapping views of the same bytes) must take the
    message-per-lane scan instead and give the same answers."""
    rng = np.random.default_rng(5)
    msgs = [w['msg'] for w in golden['whole']]
    ar = scatter_pack(eng, msgs, rng)
    rev = Arena(ar.data, ar.off[::-1].copy(), ar.len[::-1].copy())
    g, o = eng.parse(rev), O.orc_parse(rev)
    assert_parse_equal(rev, g, o, 'reversed index')
    assert_frame_equal(rev, eng.frame(rev, g, reuse_arena=True), O.orc_frame(rev, o), 'reversed index frame')
    # every message listed twice in a row: views overlap, the group is not a span
    dup = Arena(ar.data, np.repeat(ar.off, 2), np.repeat(ar.len, 2))
    assert_parse_equal(dup, eng.parse(dup), O.orc_parse(dup), 'duplicated views')
    # one group in order, the next one not: both scans inside one launch
    mix_off, mix_len = ar.off.copy(), ar.len.copy()
    for g0 in range(32, ar.n - 31, 64):
        mix_off[g0:g0 + 32] = mix_off[g0:g0 + 32][::-1].copy()
        mix_len[g0:g0 + 32] = mix_len[g0:g0 + 32][::-1].copy()
    mix = Arena(ar.data, mix_off, mix_len)
    eng.ring_build(['10.0.0.%d' % (i + 1) for i in range(16)])
    pg, qg = eng.route(mix)
    po = O.orc_parse(mix)
    assert_parse_equal(mix, pg, po, 'alternating groups')


def test_fresh_engines_give_the_same_records_every_time(golden):
    """Regression guard for allocation races: a fresh ctx per iteration (so stale device memory
    would show), records and record ranges checked against the oracle each time."""
    from shellac_b200.engine import Engine
    msgs = [w['msg'] for w in golden['whole']]
    o = None
    for it in range(12):
        e = Engine(device=0, max_msgs=1 << 14, max_bytes=1 << 24, max_records=1 << 20)
        try:
            ar = e.pack(msgs)
            if o is None:
                o = O.orc_parse(ar)
            g = e.parse(ar)
            assert_parse_equal(ar, g, o, 'fresh engine %d' % it)
            n = ar.n
            base, cnt = g.hdr_base[:n].astype(np.int64), g.n_headers[:n].astype(np.int64)
            order = np.argsort(base, kind='stable')
            b, c = base[order][cnt[order] > 0], cnt[order][cnt[order] > 0]
            assert (b[1:] >= b[:-1] + c[:-1]).all(), 'record ranges overlap'
        finally:
            e.close()


def test_host_paths_small_block_and_field_by_field_agree(golden, monkeypatch):
    """Batches of up to 4 096 messages come back through one device block and one copy
    (route_small in api.cu); SHL_SMALL=0 forces the field-by-field path the larger ones take.
    Both against the oracle, including frame + body gather continuing from either."""
    from shellac_b200.engine import Engine
    msgs = [w['msg'] for w in golden['whole'][::2]][:4000]        # at most 4 096: the one-block path qualifies
    hosts = ['10.0.0.%d' % (i + 1) for i in range(16)]
    outs = []
    for small in ('1', '0'):
        monkeypatch.setenv('SHL_SMALL', small)
        e = Engine(device=0, max_msgs=1 << 14, max_bytes=1 << 28, max_records=1 << 20)
        try:
            e.ring_build(hosts)
            ar = e.pack(msgs)
            po, qo = e.route(ar)
            o = O.orc_parse(ar)
            assert_parse_equal(ar, po, o, 'SHL_SMALL=' + small)
            fo = e.frame(ar, po, reuse_arena=True)
            of = O.orc_frame(ar, o)
            assert_frame_equal(ar, fo, of, 'frame after SHL_SMALL=' + small)
            bo, ob = e.body(ar), O.orc_body(ar, o, of)
            assert (bo.body_len[:ar.n] == ob.body_len[:ar.n]).all()
            start, order = e.group(ar.n)                      # must see THIS call's nodes, wherever they live on the device
            node = qo.node[:ar.n]
            placed = node != 0xffffffff
            assert int(start[len(hosts)]) == int(placed.sum())
            assert (np.diff(start.astype(np.int64)) == np.bincount(node[placed], minlength=len(hosts))).all()
            got = order[:int(start[len(hosts)])]
            assert (node[got] == np.repeat(np.arange(len(hosts)), np.diff(start.astype(np.int64)))).all()
            outs.append((qo.node[:ar.n].copy(), qo.hash[:ar.n].copy(), qo.hist.copy()))
        finally:
            e.close()
    assert (outs[0][0] == outs[1][0]).all() and (outs[0][1] == outs[1][1]).all() and (outs[0][2] == outs[1][2]).all()


def test_golden_streams_and_gzip(eng, golden):
    for s in golden['streams']:
        data, start = s['stream'], 0
        for exp in s['msgs']:
            ar = eng.pack([data[start:]])
            po = eng.parse(ar)
            v = MessageView(ar.message(0), po, 0, eng.frame(ar, po, reuse_arena=True), text=True)
            assert not O.diff_observation(v, {k: exp[k] for k in exp if k != 'start'})
            start += v.consumed()
        assert start == len(data)
    for name, z in golden['gzip'].items():
        ar = eng.pack([z['msg']])
        po = eng.parse(ar)
        v = MessageView(ar.message(0), po, 0, eng.frame(ar, po, reuse_arena=True), text=True)
        assert v.body().getvalue() == z['plain'] and v.message_complete(), name


# ------------------------------------------------------------------------------- config 2 shape
def test_c2_corpus_parse_matches_oracle(eng):
    ar = eng.corpus_host(20130725, 0, 60000, 0)
    g = eng.parse(ar)
    o = O.orc_parse(ar, hdr_cap=1 << 21)
    assert_parse_equal(ar, g, o, 'c2 60k')
    assert (g.code() == N.S_OK).all()
    eng.set_force_scalar(True)
    try:
        g2 = eng.parse(ar)
    finally:
        eng.set_force_scalar(False)
    assert_parse_equal(ar, g2, o, 'c2 60k scalar path')


def test_large_batches_take_the_pipelined_path_and_still_match(eng):
    """>= 64 K messages go through the chunked multi-stream path (H2D / kernel / D2H overlapped,
    one record sub-arena per chunk); a batch whose offsets are not locally dense falls back to the
    single-stream path.  Both must equal the oracle."""
    n = 150000
    ar = eng.corpus_host(99, 5000, n, 0)
    o = O.orc_parse(ar, hdr_cap=n * 24)
    g = eng.parse(ar, hdr_cap=n * 24)
    assert_parse_equal(ar, g, o, 'pipelined parse 150k')
    # same messages, index order shuffled: every chunk spans the whole arena -> fallback path
    perm = np.random.default_rng(8).permutation(n)
    a_off, a_len = eng.pinned(n, np.uint64), eng.pinned(n, np.uint32)
    a_off[:] = ar.off[perm]
    a_len[:] = ar.len[perm]
    ar2 = Arena(ar.data, a_off, a_len)
    assert_parse_equal(ar2, eng.parse(ar2, hdr_cap=n * 24), O.orc_parse(ar2, hdr_cap=n * 24), 'shuffled 150k (fallback)')


def test_c2_corpus_every_alignment_and_tail(eng):
    """Messages at all 16 alignments, and truncated at every length around the chunk edges."""
    rng = np.random.default_rng(2)
    base = eng.corpus_host(7, 0, 400, 0)
    msgs = [base.message(i) for i in range(400)]
    cut = []
    for m in msgs[:60]:
        for k in (1, 2, 3, 4, 5, 17, 33):
            cut.append(m[:len(m) - k])
        for k in (496, 510, 511, 512, 513, 514, 528):
            if k < len(m):
                cut.append(m[:k])
    ar = scatter_pack(eng, msgs + cut, rng)
    assert_parse_equal(ar, eng.parse(ar), O.orc_parse(ar), 'alignment + truncation')


def test_structural_edge_cases(eng):
    """Shapes that leave the vector path: long lines, > 32 lines per chunk, whitespace around the
    header block, lines straddling 512-byte chunks, > 640 records per warp group, 64 KiB limit."""
    rng = np.random.default_rng(3)
    msgs = []
    for n in (300, 500, 511, 512, 513, 1500, 5000):                      # long URL / first line
        msgs.append(b'GET /' + b'u' * n + b' HTTP/1.1\r\nHost: a\r\nX: y\r\n\r\n')
    for n in (480, 495, 505, 1020, 3000):                               # long header value
        msgs.append(b'GET /a HTTP/1.1\r\nHost: a\r\nBig: ' + b'v' * n + b'\r\nAfter: z\r\n\r\n')
        msgs.append(b'GET /a HTTP/1.1\r\n' + b'K' * n + b': v\r\nAfter: z\r\n\r\n')
    for n in (20, 31, 32, 33, 34, 64, 65, 100, 200, 641, 700, 2000):     # many short lines
        msgs.append(b'GET /a HTTP/1.1\r\n' + b''.join(b'a%d: b\r\n' % i for i in range(n)) + b'\r\n')
        msgs.append(b'HTTP/1.1 200 OK\r\n' + b''.join(b'h: %d\r\n' % i for i in range(n)) + b'Content-Length: 3\r\n\r\nabc')
    for pad in range(0, 40):                                            # ': ' and CRLF straddling chunk edges
        msgs.append(b'GET /' + b'p' * (460 + pad) + b' HTTP/1.1\r\nHost: example.org\r\nAccept: */*\r\n\r\n')
        msgs.append(b'GET /a HTTP/1.1\r\nHost: ' + b'h' * (470 + pad) + b'\r\nA: b\r\nC: d\r\n\r\n')
    msgs += [b'GET /a HTTP/1.1\r\n \tHost: a\r\nX: y\r\n\r\n', b'GET /a HTTP/1.1\r\nHost: a\r\nX: y \t\r\n\r\n',
             b'GET /a HTTP/1.1\r\nHost: a\r\nX: \r\n\r\n', b'GET /a HTTP/1.1\r\nHost: a\r\n \r\n\r\n',
             b'GET /a HTTP/1.1\r\nHost:  a  \r\nX:\ty\r\n\r\n', b'GET /a HTTP/1.1\r\nHost: a: b\r\n\r\n',
             b'GET /a HTTP/1.1\r\nnocolon\r\nHost: a\r\n\r\n', b'GET /a HTTP/1.1\r\nHost: a\r\nnocolon\r\n',
             b'POST /a HTTP/1.1\r\nContent-Length: 5\r\nContent-Length: 5\r\n\r\nhello',
             b'POST /a HTTP/1.1\r\ncontent-LENGTH:   5  \r\n\r\nhello', b'HEAD /a HTTP/1.0\r\nA: b\r\n\r\n',
             b'GET /a HTTP/1.1\r\nTransfer-Encoding: chunked\r\n\r\n', b'PUT /a HTTP/1.1\r\nTRANSFER-ENCODING: chunked\r\n\r\n0\r\n\r\n']
    big = b'GET /a HTTP/1.1\r\n' + b''.join(b'Header-%05d: %s\r\n' % (i, b'v' * 40) for i in range(1165)) + b'\r\n'
    assert 65000 < len(big) < 65536
    msgs += [big, big[:-2] + b'X: ' + b'y' * 400 + b'\r\n\r\n']              # just under / over the 64 KiB limit
    msgs += [m + b'GET /next HTTP/1.1\r\nA: b\r\n\r\n' for m in msgs[:20]]   # pipelined tails
    order = rng.permutation(len(msgs))
    ar = scatter_pack(eng, [msgs[i] for i in order], rng)
    g, o = eng.parse(ar), O.orc_parse(ar)
    assert_parse_equal(ar, g, o, 'structural edge cases')
    assert (g.code() == N.E_LIB_HDR_TOO_LARGE).sum() == 1
    assert_frame_equal(ar, eng.frame(ar, g, reuse_arena=True), O.orc_frame(ar, o), 'edge frame')


def test_token_soup_through_the_span_scan(eng):
    """Groups of small messages are scanned as one span of bytes: random sequences of the tokens
    the grammar is made of (so that CRLFs, ': ', spaces and their halves land next to message
    boundaries, list slices and row edges in every combination), tightly packed and with gaps."""
    rng = np.random.default_rng(11)
    toks = [b'\r\n', b'\r\n', b': ', b': ', b' ', b':', b'\r', b'\n', b'\t', b'a', b'bc', b'Host', b'GET', b'POST', b'/x/y',
            b'HTTP/1.1', b'HTTP/1.0', b'HTTP/2', b'200', b'OK', b'Content-Length', b'content-length', b'5', b'-1', b'1e3',
            b'Transfer-Encoding', b'chunked', b'\r\n\r\n', b'\x00', b'\xff', b'\x0b']
    msgs = []
    for i in range(6000):
        k = int(rng.integers(0, 40))
        m = b''.join(toks[j] for j in rng.integers(0, len(toks), k))
        if i % 3 == 0:                                    # a well-formed skeleton with soup in the values
            hdrs = b''.join(b'K%d: %s\r\n' % (h, b''.join(toks[j] for j in rng.integers(2, len(toks), int(rng.integers(0, 4)))).replace(b'\r\n', b'_'))
                            for h in range(int(rng.integers(0, 12))))
            m = b'GET /' + m.replace(b' ', b'').replace(b'\r', b'').replace(b'\n', b'')[:40] + b' HTTP/1.1\r\n' + hdrs + b'\r\n'
        msgs.append(m)
    for name, ar in (('tight', eng.pack(msgs)), ('gaps', scatter_pack(eng, msgs, rng))):
        g, o = eng.parse(ar), O.orc_parse(ar)
        assert_parse_equal(ar, g, o, 'token soup ' + name)
        assert_frame_equal(ar, eng.frame(ar, g, reuse_arena=True), O.orc_frame(ar, o), 'token soup frame ' + name)
    # every outcome the tokeniser knows shows up in this corpus
    seen = set(int(c) for c in np.unique(o.status[:ar.n] & 0xff))
    assert {N.S_OK, N.S_INCOMPLETE_LINE, N.S_INCOMPLETE_HEADERS, N.E_FIRSTLINE_PARTS, N.E_HDR_NO_SEP, N.E_HDR_MULTI_SEP} <= seen, seen


def test_empty_and_tiny_batches(eng):
    for msgs in ([b''], [b'', b'', b''], [b'G'], [b'\r\n'], [b'GET / HTTP/1.1\r\nA: b\r\n\r\n'] * 33):
        ar = eng.pack(msgs)
        assert_parse_equal(ar, eng.parse(ar), O.orc_parse(ar), repr(msgs[:1]))


# ------------------------------------------------------------------------------- config 3 shape
def _response_corpus(rng, n, lo=4096, hi=1 << 20):
    msgs = []
    exts = [b'', b'', b'', b'', b'', b'', b'', b'', b'', b';ext=foo', b';ext="foo"', b';ext=foo7']
    for i in range(n):
        size = int(np.exp(rng.uniform(np.log(lo), np.log(hi))))
        body = rng.integers(0, 256, size, dtype=np.uint8).tobytes()
        first = [b'HTTP/1.1 200 OK\r\n', b'HTTP/1.1 302 Not Modified\r\n', b'HTTP/1.1 500 Internal Server Error\r\n'][i % 3]
        hdrs = b''.join(b'X-H%d: %s\r\n' % (k, b'v' * int(rng.integers(1, 40))) for k in range(int(rng.integers(3, 11))))
        if i % 2 == 0:
            msgs.append(first + hdrs + b'Content-Length: %d\r\n\r\n' % size + body)
        else:
            out, p = [], 0
            while p < size:
                c = min(int(rng.integers(512, 16385)), size - p)
                hx = (b'%x' if rng.random() < 0.5 else b'%X') % c
                out.append(hx + exts[int(rng.integers(0, len(exts)))] + b'\r\n' + body[p:p + c] + b'\r\n')
                p += c
            out.append(b'0\r\n\r\n')
            msgs.append(first + hdrs + b'Transfer-Encoding: chunked\r\n\r\n' + b''.join(out))
    return msgs


def test_c3_responses_framing(eng):
    """Config 3: Content-Length and chunked bodies, 4 KiB - 1 MiB; complete and truncated."""
    rng = np.random.default_rng(20130726)
    msgs = _response_corpus(rng, 96)
    msgs += [m[:int(rng.integers(1, len(m)))] for m in msgs[:48]]            # truncated anywhere
    ar = scatter_pack(eng, msgs, rng)
    g, o = eng.parse(ar), O.orc_parse(ar)
    assert_parse_equal(ar, g, o, 'c3 parse')
    gf, of = eng.frame(ar, g, reuse_arena=True), O.orc_frame(ar, o)
    assert_frame_equal(ar, gf, of, 'c3 frame')
    # round trip: de-chunked body == the bytes that were chunked (size-independent property)
    for i in range(0, 96, 7):
        v = MessageView(ar.message(i), g, i, gf)
        assert v.message_complete()
        assert len(v.raw_body()) == (int(gf.body_bytes[i]) if v.chunked() else v.content_len())


def test_body_gather_matches_oracle(eng, golden):
    """N1: the de-chunk / body gather (HttpParser._parse_body) against the oracle, byte for byte,
    on the reference corpus (every alignment) and on config-3-shaped responses."""
    rng = np.random.default_rng(20130731)
    big = _response_corpus(rng, 40, lo=2000, hi=300000)
    big += [m[:int(rng.integers(1, len(m)))] for m in big[:12]]
    for name, msgs in (('golden', [w['msg'] for w in golden['whole']]), ('c3', big)):
        ar = scatter_pack(eng, msgs, rng)
        g = eng.parse(ar)
        gf = eng.frame(ar, g, reuse_arena=True)
        gb = eng.body(ar)
        o = O.orc_parse(ar)
        of = O.orc_frame(ar, o)
        ob = O.orc_body(ar, o, of)
        n = ar.n
        assert (gb.body_len[:n] == ob.body_len[:n]).all(), name
        assert int(gb.body_used[0]) == int(ob.body_used[0]), name
        assert (gb.body_off[:n] % 16 == 0).all()
        for i in range(n):
            assert gb.get(i) == ob.get(i), (name, i, ar.message(i)[:120])
        # bodies do not overlap
        nz = gb.body_len[:n] > 0
        order = np.argsort(gb.body_off[:n][nz])
        b, l = gb.body_off[:n][nz][order].astype(np.int64), gb.body_len[:n][nz][order].astype(np.int64)
        assert (b[1:] >= b[:-1] + l[:-1]).all(), name


def test_chunk_header_primitive(eng):
    cases = [b'A;ext=foo\r\nAAAA', b'\r\n8;ext="foo"\r\nB', b'0\r\n\r\n', b'zz\r\n', b'0x1f\r\n', b'5', b' 5\r\n', b'',
             b'ffffffffffffffffffff\r\n', b'1;a=b;c=d\r\n', b'1;a="b\r\n', b'1;a=b"\r\n', b'1;a=""\r\n', b'\r\n\r\n5\r\n']
    ar = eng.pack(cases)
    hl, sz, ok = eng.chunk_header(ar)
    ohl, osz, ook = O.orc_chunk_header(ar)
    assert hl.tolist() == ohl.tolist() and sz.tolist() == osz.tolist() and ok.tolist() == ook.tolist()


# ------------------------------------------------------------------------------- placement
def test_md5_and_placement_golden(eng, placement_golden):
    keys = [k.encode('latin-1') for k in placement_golden['keys']]
    rng = np.random.default_rng(4)
    ar = scatter_pack(eng, keys, rng)
    for name, r in placement_golden['rings'].items():
        eng.ring_build(r['hosts'], r['ports'], r['weights'], r['mode'])
        vals, nodes = eng.ring_export()
        ov, on = O.orc_ring(r['hosts'], r['ports'], r['weights'], r['mode'])
        assert (vals == ov).all() and (nodes == on).all() and len(vals) == r['n_points'], name
        out = eng.place(ar)
        assert [out.digest(i).hex() for i in range(len(keys))] == placement_golden['md5'], name
        assert out.hash[:len(keys)].tolist() == r['hash'], name
        assert out.node[:len(keys)].tolist() == r['node'], name
        assert out.hist.tolist() == np.bincount(r['node'], minlength=len(r['hosts'])).tolist(), name


def test_md5_all_lengths_and_alignments(eng):
    rng = np.random.default_rng(5)
    keys = [rng.integers(0, 256, n, dtype=np.uint8).tobytes() for n in list(range(0, 200)) + [250, 251, 255, 256, 257, 1000, 4096]]
    eng.ring_build(['10.0.0.%d' % (i + 1) for i in range(16)])
    for rep in range(4):
        ar = scatter_pack(eng, keys, rng)
        out = eng.place(ar)
        for i, k in enumerate(keys):
            assert out.digest(i) == hashlib.md5(k).digest(), (rep, len(k))


def test_route_fused_matches_parse_then_place(eng, golden):
    """shl_route_batch == shl_parse_batch + (key = url()) + shl_place_batch == oracle."""
    hosts = ['10.0.0.%d' % (i + 1) for i in range(16)]
    eng.ring_build(hosts)
    vals, nodes = O.orc_ring(hosts, [11211] * 16)
    rng = np.random.default_rng(6)
    c2 = eng.corpus_host(20130725, 1000, 5000, 0)
    msgs = [c2.message(i) for i in range(5000)] + [w['msg'] for w in golden['whole']]
    ar = scatter_pack(eng, msgs, rng)
    n = ar.n
    pg, qg = eng.route(ar)
    o = O.orc_parse(ar)
    assert_parse_equal(ar, pg, o, 'route parse fields')
    st = o.status[:n]
    routable = ((st & N.F_MESSAGE_COMPLETE) != 0) & ((st & N.F_IS_REQUEST) != 0) & ((st & N.F_FIRST_LINE_DONE) != 0) & ((st & 0xff) == 0) & \
               ((st & N.F_KEY_TOO_LONG) == 0)
    assert routable.sum() > 5000
    koff = (ar.off + o.sp1[:n] + 1).astype(np.uint64)
    klen = (o.sp2[:n] - o.sp1[:n] - 1).astype(np.uint32)
    koff[~routable] = 0
    klen[~routable] = 0
    oq = O.orc_place(ar, vals, nodes, 16, np.ascontiguousarray(koff), np.ascontiguousarray(klen))
    r = routable
    assert (qg.hash[:n][r] == oq.hash[:n][r]).all()
    assert (qg.node[:n][r] == oq.node[:n][r]).all()
    assert (qg.md5[:16 * n].reshape(n, 16)[r] == oq.md5[:16 * n].reshape(n, 16)[r]).all()
    assert (qg.node[:n][~r] == 0xffffffff).all() and (qg.hash[:n][~r] == 0).all()
    assert qg.hist.tolist() == np.bincount(oq.node[:n][r], minlength=16).tolist()
    # the separate entry point agrees with the fused one
    ks = eng.place(ar, np.ascontiguousarray(koff[r]), np.ascontiguousarray(klen[r]))
    assert (ks.node[:int(r.sum())] == qg.node[:n][r]).all()


def test_group_by_node_after_route(eng, golden):
    """N3: the per-node grouping pylibmc does behind get/set (Server.py:335,432)."""
    for hosts in (['10.0.0.%d' % (i + 1) for i in range(16)], ['10.3.%d.%d' % (i // 250, i % 250 + 1) for i in range(1500)]):
        eng.ring_build(hosts)
        c2 = eng.corpus_host(20130725, 7000, 20000, 0)
        msgs = [c2.message(i) for i in range(20000)] + [w['msg'] for w in golden['whole'][:1500]]
        ar = eng.pack(msgs)
        pg, qg = eng.route(ar)
        start, order = eng.group(ar.n)
        node = qg.node[:ar.n]
        placed = node != 0xffffffff
        assert int(start[len(hosts)]) == int(placed.sum()) and int(start[0]) == 0
        assert (np.diff(start.astype(np.int64)) == np.bincount(node[placed], minlength=len(hosts))).all()
        got = order[:int(start[len(hosts)])]
        assert len(np.unique(got)) == len(got) and placed[got].all()
        owner = np.repeat(np.arange(len(hosts)), np.diff(start.astype(np.int64)))
        assert (node[got] == owner).all()


def test_c2_full_size_properties(eng):
    """BASELINE config 2 at full size (1 M GETs, 16x100 ring): size-independent properties plus a
    full comparison with the oracle (the C oracle finishes 1 M messages in seconds)."""
    n = 1 << 20
    hosts = ['10.0.0.%d' % (i + 1) for i in range(16)]
    eng.ring_build(hosts)
    ar = eng.corpus_host(20130725, 0, n, 0)
    pg, qg = eng.route(ar, hdr_cap=1 << 25)
    assert (pg.code() == N.S_OK).all()
    assert int(qg.hist.sum()) == n                                    # every request placed exactly once
    assert (pg.first_line_len[:n] + pg.headers_len[:n] == ar.len).all()
    assert int(pg.n_headers[:n].sum()) == int(pg.hdr_used[0])
    hist = eng.code_hist()
    assert int(hist[N.S_OK]) >= n
    o = O.orc_parse(ar, hdr_cap=1 << 25)
    assert_parse_equal(ar, pg, o, 'c2 1M')
    vals, nodes = O.orc_ring(hosts, [11211] * 16)
    koff = (ar.off + o.sp1[:n] + 1).astype(np.uint64)
    klen = (o.sp2[:n] - o.sp1[:n] - 1).astype(np.uint32)
    oq = O.orc_place(ar, vals, nodes, 16, koff, klen)
    assert (qg.node[:n] == oq.node[:n]).all() and (qg.hash[:n] == oq.hash[:n]).all()
    assert (qg.md5[:16 * n] == oq.md5[:16 * n]).all()
    assert qg.hist.tolist() == oq.hist.tolist()


# ------------------------------------------------------------------------------- configs 4 and 5
def _host_corpus(lib, seed, first, n, kind):
    ln = np.zeros(n, np.uint32)
    lib.shl_corpus_lengths_host(seed, first, n, kind, ln.ctypes.data)
    off = np.zeros(n, np.uint64)
    np.cumsum(ln[:-1], dtype=np.uint64, out=off[1:])
    data = np.zeros(int(off[-1] + ln[-1]) + 64, np.uint8)
    lib.shl_corpus_fill_host(seed, first, n, kind, off.ctypes.data, data.ctypes.data)
    return Arena(data, off, ln)


def test_c4_device_corpus_batches_weighted_ring(eng):
    """Config 4 shape at reduced size: requests generated IN HBM batch by batch (as a rank of the
    100 M-request job does), 256-node x 160-point weighted ring, histogram accumulated on the
    device across batches; checked against the oracle over the host copy of the same corpus."""
    hosts = ['10.1.%d.%d' % (i // 250, i % 250 + 1) for i in range(256)]
    eng.ring_build(hosts, mode=N.RING_KETAMA_WEIGHTED)
    vals, nodes = eng.ring_export()
    ov, on = O.orc_ring(hosts, [11211] * 256, [1] * 256, 1)
    assert len(vals) == 40960 and (vals == ov).all() and (nodes == on).all()
    seed, per, nb = 20130727, 1 << 17, 4
    hist = np.zeros(256, np.uint64)
    ohist = np.zeros(256, np.uint64)
    for b in range(nb):
        db = eng.corpus_dev(seed, b * per, per, 0)
        har = _host_corpus(eng.lib, seed, b * per, per, 0)
        if b == 0:                                             # device bytes == host bytes for the same (seed, i)
            dev_bytes = db.fetch(db.d_arena, np.uint8, db.nbytes)
            assert db.nbytes == har.nbytes and (dev_bytes == har.data[:har.nbytes]).all()
            assert (db.len == har.len).all()
        db.route(256)
        eng.sync()
        st = db.field('status')
        assert ((st & 0xff) == 0).all()
        node = db.fetch(db.d_node, np.uint32, per)
        hist += db.fetch(db.d_hist, np.uint64, 256)
        o = O.orc_parse(har, hdr_cap=per * 24)
        oq = O.orc_place(har, ov, on, 256, (har.off + o.sp1[:per] + 1).astype(np.uint64), (o.sp2[:per] - o.sp1[:per] - 1).astype(np.uint32))
        assert (node == oq.node[:per]).all()
        assert (db.field('sp1') == o.sp1[:per]).all() and (db.field('n_headers') == o.n_headers[:per]).all()
        ohist += oq.hist
        db.free()
    assert hist.tolist() == ohist.tolist() and int(hist.sum()) == per * nb


def test_c5_remap_sweep_1024_nodes(eng):
    """Config 5 at reduced size: keys-only corpus, 1024-node ring (102 400 points); remove host
    0 / 511 / 1023 and append a new one, rebuild, re-place; node IDENTITY (host string) must equal
    the oracle's for every key, and about 1/1024 of the keys move."""
    n = 1 << 19
    base_hosts = ['10.%d.%d.%d' % (2 + i // 62500, (i // 250) % 250, i % 250 + 1) for i in range(1024)]
    db = eng.corpus_dev(20130728, 0, n, 1)
    har = _host_corpus(eng.lib, 20130728, 0, n, 1)
    assert (db.fetch(db.d_arena, np.uint8, db.nbytes) == har.data[:har.nbytes]).all()

    def place(hosts):
        eng.ring_build(hosts)
        ov, on = O.orc_ring(hosts, [11211] * len(hosts))
        v, nd = eng.ring_export()
        assert (v == ov).all() and (nd == on).all()
        db.q = None                                           # fresh histogram / node arrays per ring size
        db.place(len(hosts))
        eng.sync()
        node = db.fetch(db.d_node, np.uint32, n)
        oq = O.orc_place(har, ov, on, len(hosts))
        assert (node == oq.node[:n]).all()
        assert db.fetch(db.d_hist, np.uint64, len(hosts)).tolist() == oq.hist.tolist()
        return np.array(hosts, dtype=object)[node]

    owner0 = place(base_hosts)
    assert len(eng.ring_export()[0]) == 102400
    for k in (0, 511, 1023):
        hosts = base_hosts[:k] + base_hosts[k + 1:] + ['10.9.9.%d' % (k % 250 + 1)]
        owner = place(hosts)
        moved = owner != owner0
        was_on_removed = owner0 == base_hosts[k]
        assert moved[was_on_removed].all()                     # every key of the removed host moves
        frac = float(moved.mean())
        assert 0.5 / 1024 < frac < 4.0 / 1024, frac            # removed host's share + the new host's share
        assert set(owner[moved & ~was_on_removed]) <= {hosts[-1]}   # anything else that moved went to the new host
    db.free()


def test_header_names_lowercased_on_the_device(eng, golden):
    """R3, key.lower() (HttpParser.py:330): the names come back lower-cased FROM THE KERNEL (no host
    .lower() anywhere in this test) and equal the keys of the reference's own headers dict."""
    from shellac_b200.batch import ParseOut
    whole = golden['whole']
    msgs = [w['msg'] for w in whole]
    ar = eng.pack(msgs)
    n = ar.n
    out = ParseOut(n, ar.nbytes // 4 + 64, alloc=eng.pinned, lkeys_cap=ar.nbytes + 64)
    pg = eng.parse(ar, out=out)
    o = O.orc_parse(ar, lkeys=True)
    assert int(pg.lkeys_used[0]) == int(o.lkeys_used[0]) > 10000
    checked = mixed = 0
    for i, w in enumerate(whole):
        got = pg.lower_keys(i)
        assert got == o.lower_keys(i), w['id']
        if w['obs']['exc'] is None and w['obs']['headers_complete']:
            assert set(k.decode('latin-1') for k in got) == set(w['obs']['headers'].keys()), w['id']
            checked += 1
            b, k = int(pg.hdr_base[i]), int(pg.n_headers[i])
            raw = [ar.message(i)[int(r['key_off']):int(r['key_off']) + int(r['key_len'])] for r in pg.hdr[b:b + k]]
            mixed += sum(x != y for x, y in zip(raw, got))
    assert checked > 2500 and mixed > 1000          # plenty of names really had upper-case letters
    # the fused route entry fills the same arena
    eng.ring_build(['10.0.0.%d' % (i + 1) for i in range(16)])
    out2 = ParseOut(n, ar.nbytes // 4 + 64, alloc=eng.pinned, lkeys_cap=ar.nbytes + 64)
    p2, _ = eng.route(ar, pout=out2)
    assert all(p2.lower_keys(i) == pg.lower_keys(i) for i in range(0, n, 7))


def test_key_too_long_is_flagged_not_placed(eng):
    """R9: the cache key is the URL (Server.py:327); pylibmc refuses keys over 250 bytes before it
    hashes them (Server.py:335).  Such requests parse normally, carry SHL_F_KEY_TOO_LONG and get no node."""
    hosts = ['10.0.0.%d' % (i + 1) for i in range(16)]
    eng.ring_build(hosts)
    lens = [1, 249, 250, 251, 252, 400, 3000]
    msgs = [b'GET /' + b'k' * (L - 1) + b' HTTP/1.1\r\nHost: x\r\n\r\n' for L in lens]
    msgs += [b'GET /a b' + b'k' * 300 + b' HTTP/1.1\r\nHost: x\r\n\r\n']        # the key is "/a": short
    ar = eng.pack(msgs * 40)
    pg, qg = eng.route(ar)
    o = O.orc_parse(ar)
    assert (pg.status[:ar.n] == o.status[:ar.n]).all()
    flag = (pg.status[:ar.n] & N.F_KEY_TOO_LONG) != 0
    want = np.array([L > N.MAX_KEY_LEN for L in lens] + [False] * 1, bool)
    assert (flag == np.tile(want, 40)).all()
    assert (qg.node[:ar.n][flag] == 0xffffffff).all() and (qg.node[:ar.n][~flag] < 16).all()
    assert int(qg.hist.sum()) == int((~flag).sum())
    for i in np.flatnonzero(~flag)[:8]:
        key = ar.message(int(i))[int(pg.sp1[i]) + 1:int(pg.sp2[i])]
        assert qg.digest(int(i)) == hashlib.md5(key).digest()


def test_memcached_frames_per_node(eng, golden):
    """N3: get / set request frames of the memcached binary protocol, one contiguous stream per node
    (what pylibmc sends behind Server.py:335 / :432), against the struct.pack restatement."""
    import os, sys
    sys.path.insert(0, os.path.join(O.ROOT, 'oracle'))
    import py_port
    hosts = ['10.0.0.%d' % (i + 1) for i in range(16)]
    eng.ring_build(hosts)
    c2 = eng.corpus_host(20130725, 31000, 6000, 0)
    msgs = [c2.message(i) for i in range(6000)] + [w['msg'] for w in golden['whole'][:1500]]
    ar = eng.pack(msgs)
    n = ar.n
    pg, qg = eng.route(ar)
    start, order = eng.group(n)
    n_placed = int(start[len(hosts)])
    assert n_placed > 6000
    foff, frames = eng.mc_frames(n, n_placed, N.MC_GET)
    blob = frames.tobytes()
    assert int(foff[n_placed]) == len(blob)
    for k in range(len(hosts)):                                  # node k's stream: its requests' frames back to back
        want = b''
        for j in range(int(start[k]), int(start[k + 1])):
            i = int(order[j])
            key = ar.message(i)[int(pg.sp1[i]) + 1:int(pg.sp2[i])]
            assert int(qg.node[i]) == k
            want += py_port.mc_get_frame(key, i)
        assert blob[int(foff[start[k]]):int(foff[start[k + 1]])] == want, k
    vlen = (np.arange(n, dtype=np.uint32) * 37) % 5000
    foff, frames = eng.mc_frames(n, n_placed, N.MC_SET, ttl=170, value_len=vlen)       # ttl default of the reference CLI (Server.py:510)
    blob = frames.tobytes()
    for j in list(range(0, n_placed, 97)) + [n_placed - 1]:
        i = int(order[j])
        key = ar.message(i)[int(pg.sp1[i]) + 1:int(pg.sp2[i])]
        assert blob[int(foff[j]):int(foff[j + 1])] == py_port.mc_set_header(key, i, int(vlen[i]), 170), j


def test_compact_records_equal_the_wide_layout(eng, golden):
    """shl_route_compact_batch: the 32-byte record per message says what the struct-of-arrays fields and
    the node of shl_route_batch say (small batch and the chunked large-batch path)."""
    hosts = ['10.0.0.%d' % (i + 1) for i in range(16)]
    eng.ring_build(hosts)
    c2 = eng.corpus_host(20130725, 50000, 70000, 0)
    for msgs in ([w['msg'] for w in golden['whole']], [c2.message(i) for i in range(70000)] + [w['msg'] for w in golden['whole'][:500]]):
        ar = eng.pack(msgs)
        n = ar.n
        pg, qg = eng.route(ar)
        rec, hdr, used, md5, hist = eng.route_compact(ar, want_md5=True)
        rec = rec[:n]
        wide_first = pg.first_line_len[:n] > 65535
        st = pg.status[:n].copy()
        assert not wide_first.any()
        sc = pg.status_code[:n].astype(np.int64)
        resp_wide = ((st & N.F_IS_REQUEST) == 0) & ((sc > 65535) | (sc < 0))
        st[resp_wide] |= N.F_STATUS_WIDE
        assert (rec['status'] == st).all()
        for f in ('first_line_len', 'headers_len', 'sp1', 'sp2', 'fl_end', 'slash', 'n_headers', 'content_len'):
            assert (rec[f].astype(np.int64) == getattr(pg, f)[:n].astype(np.int64)).all(), f
        req = (st & N.F_IS_REQUEST) != 0
        assert (rec['tail'][req] == np.minimum(qg.node[:n][req], 0xffff)).all()
        ok_resp = ~req & ((st & N.F_STATUS_WIDE) == 0)
        assert (rec['tail'][ok_resp] == (pg.status_code[:n][ok_resp].astype(np.int64) & 0xffff)).all()
        assert (md5[:16 * n] == qg.md5[:16 * n]).all() and hist[:16].tolist() == qg.hist.tolist()
        assert int(used[0]) == int(pg.hdr_used[0])
        # same header records behind hdr_base
        a = _records(pg, n)
        class R: pass
        r = R(); r.n_headers, r.hdr_base, r.hdr = rec['n_headers'].astype(np.uint32), rec['hdr_base'], hdr
        assert (_records(r, n) == a).all()


def test_mutated_corpus_messages(eng):
    """Well-formed config-2 requests with a few bytes overwritten, inserted or deleted (CR, LF, ':', ' ',
    tab, NUL, a letter): near-valid structure is what the bitmap path's alternation check, first-line
    shortcut and value strip have to get right.  route (fused) vs the oracle, 3 seeds x 20 000 messages."""
    hosts = ['10.0.0.%d' % (i + 1) for i in range(16)]
    eng.ring_build(hosts)
    vals, nodes = O.orc_ring(hosts, [11211] * 16)
    c2 = eng.corpus_host(20130725, 200000, 20000, 0)
    base = [c2.message(i) for i in range(20000)]
    pool = [b'\r', b'\n', b':', b' ', b'\t', b'\x00', b'a', b'\r\n', b': ', b'  ', b'\r\n\r\n', b'']
    seen = set()
    for seed in (1, 2, 3):
        rng = np.random.default_rng(seed)
        msgs = []
        for m in base:
            b = bytearray(m)
            for _ in range(int(rng.integers(0, 4))):
                p = int(rng.integers(0, len(b) + 1))
                t = pool[int(rng.integers(0, len(pool)))]
                op = int(rng.integers(0, 3))
                if op == 0:
                    b[p:p + len(t)] = t                 # overwrite
                elif op == 1:
                    b[p:p] = t                          # insert
                else:
                    del b[p:p + int(rng.integers(1, 4))]   # delete
            msgs.append(bytes(b))
        ar = eng.pack(msgs)
        n = ar.n
        pg, qg = eng.route(ar)
        o = O.orc_parse(ar)
        assert_parse_equal(ar, pg, o, 'mutated seed %d' % seed)
        st = o.status[:n]
        r = ((st & N.F_MESSAGE_COMPLETE) != 0) & ((st & N.F_IS_REQUEST) != 0) & ((st & N.F_FIRST_LINE_DONE) != 0) & \
            ((st & 0xff) == 0) & ((st & N.F_KEY_TOO_LONG) == 0)
        koff = (ar.off + o.sp1[:n] + 1).astype(np.uint64); klen = (o.sp2[:n] - o.sp1[:n] - 1).astype(np.uint32)
        koff[~r] = 0; klen[~r] = 0
        oq = O.orc_place(ar, vals, nodes, 16, np.ascontiguousarray(koff), np.ascontiguousarray(klen))
        assert (qg.node[:n][r] == oq.node[:n][r]).all() and (qg.node[:n][~r] == 0xffffffff).all()
        seen |= set(int(c) for c in np.unique(st & 0xff))
    assert {N.S_OK, N.S_INCOMPLETE_HEADERS, N.E_FIRSTLINE_PARTS, N.E_HDR_NO_SEP, N.E_HDR_MULTI_SEP} <= seen, seen
