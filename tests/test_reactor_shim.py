"""N4: the reactor batching shim.  Host logic on CPU with an oracle-backed router, and the same
scenario through the real engine with -m gpu."""
import numpy as np
import pytest

from shellac_b200.batch import Arena
from shellac_b200.reactor import ReactorShim
from tests import oracle_lib as O

HOSTS = ['10.0.0.%d' % (i + 1) for i in range(16)]


class _Bag(object):
    pass


class OracleRouter(object):
    """TEST-ONLY: answers route() from the CPU oracle (lives in tests/, never in the package)."""

    def __init__(self):
        self.vals, self.nodes = O.orc_ring(HOSTS, [11211] * len(HOSTS))

    def route(self, messages):
        ar = Arena.pack(messages)
        n = ar.n
        po = O.orc_parse(ar)
        fo = O.orc_frame(ar, po)
        st = po.status[:n]
        ok = ((st & 0x400) != 0) & ((st & 0x100) != 0) & ((st & (1 << 17)) != 0) & ((st & 0xff) == 0) & ((st & (1 << 18)) == 0)
        koff = (ar.off + po.sp1[:n] + 1).astype(np.uint64)
        klen = (po.sp2[:n] - po.sp1[:n] - 1).astype(np.uint32)
        koff[~ok] = 0
        klen[~ok] = 0
        q = O.orc_place(ar, self.vals, self.nodes, len(HOSTS), np.ascontiguousarray(koff), np.ascontiguousarray(klen))
        q.node[:n][~ok] = 0xffffffff
        return po, q, fo

    def place_keys(self, keys):
        ar = Arena.pack(keys)
        return O.orc_place(ar, self.vals, self.nodes, len(HOSTS)).node[:len(keys)]


def _scenario(router, golden):
    import random
    rng = random.Random(5)
    shim = ReactorShim(router)
    streams = {s['id']: s for s in golden['streams']}
    t3 = streams['T3_request_stream']['stream']                     # GET, POST(10), POST(20) pipelined
    chunked_put = ('PUT /up HTTP/1.1\r\nTransfer-Encoding: chunked\r\n\r\n5\r\nhello\r\n0\r\n\r\n'
                   'GET /after HTTP/1.1\r\nA: b\r\n\r\n')
    conns = {1: t3, 2: 'GET /a HTTP/1.1\r\nHost: x\r\n\r\n', 3: chunked_put, 4: 'GET /bad\r\nHost: x\r\n\r\n',
             5: 'GET /partial HTTP/1.1\r\nHost: x\r\n'}
    cursor = {fd: 0 for fd in conns}
    got = {fd: [] for fd in conns}
    for tick in range(40):
        for fd, data in conns.items():
            if cursor[fd] < len(data):
                k = rng.randint(1, 60)
                shim.feed(fd, data[cursor[fd]:cursor[fd] + k])
                cursor[fd] += k
        for r in shim.tick():
            got[r.fd].append(r)
    assert [r.key for r in got[1]] == [b'/stream1.html', b'/stream2.html', b'/stream3.html']
    assert [r.view.method() for r in got[1]] == [b'GET', b'POST', b'POST']
    assert b''.join(r.raw for r in got[1]).decode('latin-1') == t3
    assert [r.key for r in got[2]] == [b'/a']
    assert [r.key for r in got[3]] == [b'/up', b'/after'] and got[3][0].view.raw_body() == b'hello'
    assert len(got[4]) == 1 and isinstance(got[4][0].error, ValueError)
    assert got[5] == [] and bytes(shim.pending[5]).decode() == conns[5]     # incomplete: bytes stay queued
    vals, nodes = O.orc_ring(HOSTS, [11211] * len(HOSTS))
    for fd in (1, 2, 3):
        for r in got[fd]:
            ar = Arena.pack([r.key])
            assert r.node == int(O.orc_place(ar, vals, nodes, len(HOSTS)).node[0]), r.key


def test_reactor_shim_host_logic(golden):
    _scenario(OracleRouter(), golden)


@pytest.mark.gpu
def test_reactor_shim_on_device(golden):
    from shellac_b200.engine import Engine
    from shellac_b200.reactor import EngineRouter
    eng = Engine(device=0, max_msgs=1 << 12, max_bytes=1 << 22)
    try:
        eng.ring_build(HOSTS)
        _scenario(EngineRouter(eng), golden)
    finally:
        eng.close()


@pytest.mark.gpu
def test_engine_router_large_batch_keeps_every_header():
    """A batch of >= 65 536 messages takes the chunked host path, whose header records lie in per-chunk
    slices of the arena: the router's detached copy must still hold the LAST chunk's records."""
    from shellac_b200.batch import MessageView
    from shellac_b200.engine import Engine
    from shellac_b200.reactor import EngineRouter
    eng = Engine(device=0, max_msgs=1 << 17, max_bytes=1 << 26)
    eng.ring_build(HOSTS)
    c2 = eng.corpus_host(20130725, 0, 70000, 0)
    msgs = [c2.message(i) for i in range(70000)]
    pout, qout, fout = EngineRouter(eng, msgs=1 << 17, nbytes=1 << 26).route(msgs)
    o = O.orc_parse(Arena.pack(msgs))
    for i in (0, 1, 34999, 65535, 65536, 69998, 69999):
        got = MessageView(msgs[i], pout, i, fout).header_items()
        assert got == MessageView(msgs[i], o, i).header_items() and len(got) >= 8, i
    eng.close()
