"""The drop-in class surface (shellac_b200.server.HttpParser / StreamBuf) against the golden
records of the unmodified reference.  The same checks run twice: on CPU with the oracle answering
the device questions (host logic only), and with -m gpu through the real C ABI."""
import pytest

from shellac_b200.server import HttpParser, StreamBuf
from tests.oracle_backend import OracleBackend


def _observe(p, calls, exc):
    hd = {k: (list(v) if isinstance(v, list) else v) for k, v in p.headers().items()}
    ver = p.version()
    return {'calls': calls, 'exc': exc, 'is_request': p.is_request(), 'method': p.method(), 'url': p.url(),
            'version': repr(ver) if ver is not None else None, 'status': p.status(), 'message': p.message(),
            'headers': hd, 'headers_complete': p.headers_complete(), 'message_complete': p.message_complete(),
            'on_body': p._on_body, 'chunked': p._chunked, 'content_len': p._content_len,
            'body': p.body().getvalue()}


def _drive_whole(msg, be):
    p = HttpParser(backend=be)
    data, calls, exc = msg, [], None
    try:
        while data and not p.message_complete():
            c = p.parse(data, len(data))
            calls.append(c)
            data = data[c:]
    except Exception as e:          # noqa: BLE001
        exc = type(e).__name__
    return _observe(p, calls, exc)


def _drive_fragments(frags, be):
    p = HttpParser(backend=be)
    calls, exc = [], None
    try:
        for f in frags:
            req, per = f, []
            while len(req) != 0:
                c = p.parse(req, len(req))
                per.append(c)
                if c == 0:
                    break
                req = req[c:]
            calls.append(per)
    except Exception as e:          # noqa: BLE001
        exc = type(e).__name__
    return _observe(p, calls, exc)


def _check_incremental(golden, be, step):
    bad = []
    for w in golden['whole'][::step]:
        got = _drive_whole(w['msg'], be)
        d = {k: (got[k], w['obs'][k]) for k in got if got[k] != w['obs'][k]}
        if d:
            bad.append((w['id'], w['msg'][:80], d))
    for w in golden['fragments'][::max(1, step // 4)]:
        got = _drive_fragments(w['frags'], be)
        d = {k: (got[k], w['obs'][k]) for k in got if got[k] != w['obs'][k]}
        if d:
            bad.append((w['id'], w['frags'], d))
    assert not bad, (len(bad), bad[:3])


def _check_batch(golden, be):
    msgs = [w['msg'] for w in golden['whole']]
    ps = HttpParser.parse_many(msgs, backend=be)
    bad = []
    for p, w in zip(ps, golden['whole']):
        o = w['obs']
        got = {'calls': p.calls, 'exc': type(p.error).__name__ if p.error else None, 'method': p.method(), 'url': p.url(),
               'status': p.status(), 'message': p.message(), 'headers': p.headers(), 'message_complete': p.message_complete(),
               'headers_complete': p.headers_complete(), 'body': p.body().getvalue(), 'is_request': p.is_request()}
        d = {k: (got[k], o[k]) for k in got if got[k] != o[k]}
        if d:
            bad.append((w['id'], d))
    assert not bad, (len(bad), bad[:3])


def test_incremental_facade_host_logic(golden):
    _check_incremental(golden, OracleBackend(), step=1)


def test_batch_facade_host_logic(golden):
    _check_batch(golden, OracleBackend())


def test_t8_large_body_in_pieces(golden):
    """test/HttpParserTests.py:257-290 with the seeded stand-in blob."""
    import random
    t8 = golden['t8']
    rng = random.Random(t8['blob_unit_seed'])
    valch = [chr(c) for c in range(0x21, 0x7f) if chr(c) not in ':_']
    blob = ''.join(rng.choice(valch) for _ in range(4096)) * t8['blob_repeat'] + t8['blob_tail']
    req = 'HTTP/1.1 200 OK\r\nUser-Agent: IE 6\r\nContent-Length: %d\r\n\r\n' % len(blob) + blob
    pieces = [req[i:i + t8['piece']] for i in range(0, len(req), t8['piece'])]
    got = _drive_fragments(pieces, OracleBackend())
    assert got['message_complete'] and len(got['body']) == t8['obs']['body_len'] == len(blob)
    assert got['calls'] == t8['obs']['calls']


def test_reference_usage_example_and_str():
    """The usage block in the reference's docstring (HttpParser.py:5-13) and __str__ (:111-138)."""
    be = OracleBackend()
    buf = b'HTTP/1.1 500 Internal Server Error\r\nServer: Apache 2.2\r\nDate: Never\r\nContent-Length: 12\r\n\r\nRRRRRRRRRRRR'
    p = HttpParser(backend=be)
    while not p.message_complete():
        c = p.parse(buf, len(buf))
        buf = buf[c:]
    assert p.headers()[b'server'] == b'Apache 2.2' and p.body().read() == b'RRRRRRRRRRRR'
    s = str(p)
    assert s.startswith('HTTP/1.1 500 Internal Server Error\r\n') and s.endswith('\r\n\r\nRRRRRRRRRRRR')
    assert 'Server: Apache 2.2\r\n' in s and 'Content-Length: 12\r\n' in s and 'Date: Never\r\n' in s
    q = HttpParser(backend=be)
    req = 'GET /index.html HTTP/1.1\r\nUser-Agent: Mozilla/WebKit 2.11\r\nDate: Never\r\n\r\n'
    while not q.message_complete():
        c = q.parse(req, len(req))
        req = req[c:]
    assert str(q) == 'GET /index.html HTTP/1.1\r\nUser-Agent: Mozilla/WebKit 2.11\r\nDate: Never\r\n\r\n'
    assert q.keep_alive() is False and q.keep_alive_params() == (0, 1)
    r = HttpParser.parse_many([b'HTTP/1.1 200 OK\r\nConnection: Keep-Alive\r\nKeep-Alive: timeout=7, max=42\r\nContent-Length: 0\r\n\r\n'], backend=be)[0]
    assert r.keep_alive() and r.keep_alive_params() == (7, 42)


def test_streambuf_script(golden):
    """test/StreamBufTests.py:7-36 replayed; expected values recorded from the reference class."""
    s = StreamBuf()
    log = [('ready', s.ready()), ('closed', s.closed())]
    s.write('Hello'); log.append(('ready', s.ready()))
    log += [('read', s.read()), ('read', s.read())]
    s.ack(2); log.append(('read', s.read()))
    s.ack(3); log.append(('read', s.read()))
    s.close(); log.append(('closed', s.closed()))
    log.append(('buffer', s.buffer()))
    s.clear(); log += [('buffer', s.buffer()), ('ready', s.ready()), ('closed', s.closed())]
    s.write('Romeo, oh Romeo.'); s.close(); s.ack(16); log.append(('complete', s.complete()))
    assert [list(x) for x in log] == golden['streambuf']
    b = StreamBuf(b'Hello')
    b.write(b', world!')
    assert b.read() == b'Hello, world!'
    b.ack(7)
    assert b.read() == b'world!' and bytes(b.window()) == b'world!'
    b.seek(0)
    assert b.read() == b'Hello, world!' and not b.complete()


@pytest.mark.gpu
def test_incremental_facade_on_device(golden):
    from shellac_b200.server.HttpParser import DeviceBackend
    be = DeviceBackend()
    try:
        _check_incremental(golden, be, step=16)
    finally:
        be.eng.close()


@pytest.mark.gpu
def test_batch_facade_on_device(golden):
    from shellac_b200.server.HttpParser import DeviceBackend
    be = DeviceBackend()
    try:
        _check_batch(golden, be)
    finally:
        be.eng.close()
