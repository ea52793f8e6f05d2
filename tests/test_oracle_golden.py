"""The oracle (oracle/shl_oracle.c) pinned against the golden vectors the UNMODIFIED reference
produced (tests/golden/make_golden.py): every literal of test/HttpParserTests.py plus seeded
fuzz / poison / chunked corpora, and the placement restatement."""
import hashlib
import struct

import numpy as np

from shellac_b200 import _native as N
from shellac_b200.batch import Arena, MessageView
from tests import oracle_lib as O


def test_oracle_whole_messages_match_reference(golden):
    msgs = [w['msg'] for w in golden['whole']]
    ar = Arena.pack(msgs)
    po = O.orc_parse(ar)
    fo = O.orc_frame(ar, po)
    bad = []
    for i, w in enumerate(golden['whole']):
        v = MessageView(ar.message(i), po, i, fo, text=True)
        d = O.diff_observation(v, w['obs'])
        if d:
            bad.append((w['id'], w['msg'][:80], d))
    assert not bad, bad[:5]
    assert len(msgs) > 3900


def test_oracle_reference_test_literals(golden):
    """The asserts of test/HttpParserTests.py:43-46, 60-63, 212 restated on the oracle output."""
    by_id = {w['id']: w for w in golden['whole']}
    for name in ('T1_get', 'T2_post', 'T6_chunked'):
        ar = Arena.pack([by_id[name]['msg']])
        po = O.orc_parse(ar)
        v = MessageView(ar.message(0), po, 0, O.orc_frame(ar, po), text=True)
        assert v.message_complete()
        if name == 'T1_get':
            assert (v.method(), v.url(), v.version()) == ('GET', '/get-request.html', 1.1)
            assert v.headers()['user-agent'] == 'Safari'
            assert v.calls() == [32, 54]
        if name == 'T2_post':
            assert (v.method(), v.url(), v.version()) == ('POST', '/post-request.html', 1.1)
        if name == 'T6_chunked':
            assert v.body().read() == 'AAAAAAAAAABBBBBBBBCCCCCC'
            assert v.calls() == [17, 107, 21, 23, 20, 7]
            assert v.headers()['user-agent'] == ['Safari', 'Mac OS 10.8']


def test_oracle_pipelined_streams(golden):
    """test/HttpParserTests.py:85-93 and :141-149: 3 requests / 5 (+1 gzip) responses."""
    for s in golden['streams']:
        data, start, got = s['stream'], 0, 0
        for exp in s['msgs']:
            assert exp['start'] == start
            ar = Arena.pack([data[start:]])
            po = O.orc_parse(ar)
            v = MessageView(ar.message(0), po, 0, O.orc_frame(ar, po), text=True)
            d = O.diff_observation(v, {k: exp[k] for k in exp if k != 'start'})
            assert not d, (s['id'], got, d)
            start += v.consumed()
            got += 1
        assert start == len(data)
    assert [len(s['msgs']) for s in golden['streams']] == [3, 5]


def test_oracle_gzip_fixtures(golden):
    for name, z in golden['gzip'].items():
        ar = Arena.pack([z['msg']])
        po = O.orc_parse(ar)
        v = MessageView(ar.message(0), po, 0, O.orc_frame(ar, po), text=True)
        assert v.body().getvalue() == z['plain'], name
        assert len(v.calls()) == z['n_calls'] and v.message_complete()


def test_oracle_md5_rfc1321_and_hashlib(placement_golden):
    for k, hexd in placement_golden['rfc1321'].items():
        assert O.orc_md5(k.encode('latin-1')).hex() == hexd
    rng = np.random.default_rng(7)
    for n in list(range(0, 130)) + [255, 256, 1000]:
        b = rng.integers(0, 256, n, dtype=np.uint8).tobytes()
        assert O.orc_md5(b) == hashlib.md5(b).digest()


def test_oracle_placement_matches_restatement(placement_golden):
    keys = [k.encode('latin-1') for k in placement_golden['keys']]
    ar = Arena.pack(keys)
    for name, r in placement_golden['rings'].items():
        vals, nodes = O.orc_ring(r['hosts'], r['ports'], r['weights'], r['mode'])
        assert len(vals) == r['n_points'], name
        packed = struct.pack('<%dI' % (2 * len(vals)), *(list(map(int, vals)) + list(map(int, nodes))))
        assert hashlib.md5(packed).hexdigest() == r['ring_md5'], name
        out = O.orc_place(ar, vals, nodes, len(r['hosts']))
        assert out.hash[:len(keys)].tolist() == r['hash'], name
        assert out.node[:len(keys)].tolist() == r['node'], name
        assert [out.digest(i).hex() for i in range(len(keys))] == placement_golden['md5']
        assert int(out.hist.sum()) == len(keys)


def test_oracle_chunk_header_primitive():
    cases = [(b'A;ext=foo\r\nAAAA', 11, 10, 1), (b'\r\n8;ext="foo"\r\nB', 15, 8, 1), (b'0\r\n\r\n', 3, 0, 1),
             (b'zz\r\n', 4, 0, 0), (b'0x1f\r\n', 6, 31, 1), (b'5', -1, 0, 0), (b' 5\r\n', -1, 0, 0), (b'', -1, 0, 0)]
    ar = Arena.pack([c[0] for c in cases])
    hl, sz, ok = O.orc_chunk_header(ar)
    for i, (_, ehl, esz, eok) in enumerate(cases):
        assert (int(hl[i]), int(sz[i]), int(ok[i])) == (ehl, esz, eok), cases[i]


# ----------------------------------------------------------------------------- oracle/py_port.py
def _port_observation(p, calls, exc):
    def t(x):
        return x.decode('latin-1') if isinstance(x, bytes) else x
    hd = {t(k): ([t(x) for x in v] if isinstance(v, list) else t(v)) for k, v in p.headers.items()}
    return {'calls': calls, 'exc': exc, 'is_request': p.is_request, 'method': t(p.method), 'url': t(p.url),
            'version': repr(p.version) if p.version is not None else None, 'status': p.status,
            'message': t(p.reason), 'headers': hd, 'headers_complete': p.headers_complete,
            'message_complete': p.message_complete, 'chunked': p.chunked, 'content_len': p.content_len,
            'body': p.body.getvalue().decode('latin-1')}


def test_py_port_matches_reference_golden(golden):
    import sys, os
    sys.path.insert(0, os.path.join(O.ROOT, 'oracle'))
    import py_port
    bad = []
    for w in golden['whole']:
        p, calls, exc = py_port.drive(w['msg'].encode('latin-1'))
        got = _port_observation(p, calls, exc)
        d = {k: (got[k], w['obs'][k]) for k in got if got[k] != w['obs'][k]}
        if d:
            bad.append((w['id'], w['msg'][:80], d))
    assert not bad, (len(bad), bad[:3])
    for name, z in golden['gzip'].items():
        p, calls, exc = py_port.drive(z['msg'].encode('latin-1'))
        assert exc is None and p.message_complete and p.body.getvalue().decode() == z['plain']


def test_py_port_placement_matches_golden(placement_golden):
    import sys, os
    sys.path.insert(0, os.path.join(O.ROOT, 'oracle'))
    import py_port
    for name, r in placement_golden['rings'].items():
        if r['n_points'] > 5000:
            continue
        values, nodes = py_port.build_ring(r['hosts'], r['ports'], r['weights'], r['mode'])
        assert len(values) == r['n_points']
        got = [py_port.place(values, nodes, len(r['hosts']), k.encode('latin-1')) for k in placement_golden['keys']]
        assert got == r['node'], name


def test_oracle_body_gather_matches_reference_bodies(golden):
    """orc_body == what the reference's body() holds (identity content-coding cases)."""
    msgs = [w['msg'] for w in golden['whole']]
    ar = Arena.pack(msgs)
    po = O.orc_parse(ar)
    fo = O.orc_frame(ar, po)
    bo = O.orc_body(ar, po, fo)
    for i, w in enumerate(golden['whole']):
        if 'gzip' in w['obs']['headers'].get('content-encoding', ''):
            continue
        if w['obs']['exc'] and w['obs']['headers_complete'] is False:
            continue
        v = MessageView(ar.message(i), po, i, fo)
        assert bo.get(i) == v.raw_body(), w['id']
        if not w['obs']['exc']:
            assert bo.get(i).decode('latin-1') == w['obs']['body'], w['id']


def test_oracle_lower_keys_and_key_too_long(golden):
    """key.lower() (HttpParser.py:330) as the oracle emits it == the keys of the reference's headers
    dict; SHL_F_KEY_TOO_LONG marks request URLs over 250 bytes (pylibmc's limit, Server.py:335)."""
    whole = golden['whole']
    ar = Arena.pack([w['msg'] for w in whole])
    o = O.orc_parse(ar, lkeys=True)
    n_checked = 0
    for i, w in enumerate(whole):
        if w['obs']['exc'] is None and w['obs']['headers_complete']:
            assert set(k.decode('latin-1') for k in o.lower_keys(i)) == set(w['obs']['headers'].keys()), w['id']
            n_checked += 1
    assert n_checked > 2500
    msgs = [b'GET /' + b'k' * (L - 1) + b' HTTP/1.1\r\nA: b\r\n\r\n' for L in (249, 250, 251)] + [b'HTTP/1.1 200 ' + b'k' * 300 + b'\r\n\r\n']
    st = O.orc_parse(Arena.pack(msgs)).status[:4]
    assert [bool(x & N.F_KEY_TOO_LONG) for x in st] == [False, False, True, False]


def test_memcached_frame_restatement_known_answers():
    """The binary-protocol restatement against the protocol document's own examples ("Get" of key
    'Hello'; "Set"-shaped header arithmetic)."""
    import sys, os
    sys.path.insert(0, os.path.join(O.ROOT, 'oracle'))
    import py_port
    assert py_port.mc_get_frame(b'Hello', 0) == bytes.fromhex('80000005000000000000000500000000' '0000000000000000') + b'Hello'
    h = py_port.mc_set_header(b'Hello', 0xdeadbeef, 5, 0x0e10, flags=0xdeadbeef)
    # the document's "Add" example carries flags 0xdeadbeef, expiry 0x0e10, key Hello, value World: body 0x12
    assert h[2:4] == b'\x00\x05' and h[4] == 8 and h[8:12] == b'\x00\x00\x00\x12' and h[24:32] == bytes.fromhex('deadbeef00000e10')
    assert h[32:] == b'Hello' and h[12:16] == bytes.fromhex('deadbeef') and h[1] == 1
