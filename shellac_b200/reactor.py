"""Reactor batching shim: the caller-side change that lets a live,
single-threaded epoll reactor use the batch path.

The reference parses inside its read handlers, one `recv(4096)` at a time, one connection at a time
(`Server._read_requests`, Server.py:302-377: `parsed = request.parse(data, len(data))` ...
`key = request.url()` ... `self._mc.get(key)`).  With this shim the handlers only `feed(fd, data)`;
once per epoll tick `tick()` packs every connection's pending bytes into ONE arena, makes ONE
`route` call (tokenise + key + placement on the device) and hands back, per connection, the
requests that completed, in arrival order, each with its cache key and memcached node.  Bytes of
a request that is not complete yet stay queued for the next tick, exactly like the reference's
parser keeps them in `_buf` (HttpParser.py:153,174,209); pipelined requests on one connection are
peeled off in further rounds of the same tick (Server.py:320 `while len(data) != 0`).

`router` is anything with `route(messages) -> (parse_out, place_out, frame_out_or_None)`;
`EngineRouter` below is the real one.  There is no CPU path in this package.
"""
import numpy as np

from . import _native as N
from .batch import MessageView


class Routed(object):
    """One completed request: what Server._read_requests has in hand at Server.py:325-335.
    `.view` (the parser-shaped MessageView of the request) is built when it is first asked for."""
    __slots__ = ('fd', '_view', '_src', 'key', 'node', 'error', 'raw')

    def __init__(self, fd, view, key, node, error, raw, src=None):
        self.fd, self._view, self._src, self.key, self.node, self.error, self.raw = fd, view, src, key, node, error, raw

    @property
    def view(self):
        if self._view is None and self._src is not None:
            msg, pout, i, fout = self._src
            self._view = MessageView(msg, pout, i, fout)
        return self._view


class EngineRouter(object):
    """route() through an Engine: fused parse + key + placement, chunk framing when needed.

    The pinned arena and result arrays are kept between ticks and only grow (page-locking memory
    costs far more than a tick's worth of parsing), so a tick is: one join + copy of the pending
    bytes into the arena, one C call, and copies of the rows that were filled."""

    def __init__(self, engine, msgs=256, nbytes=1 << 18):
        from .engine import PinnedBatch
        self.eng = engine
        self._pb = PinnedBatch(engine, msgs, nbytes)

    def route(self, messages):
        eng = self.eng
        if not len(messages):
            ar = eng.pack(messages)
            pout, qout = eng.route(ar)
            return _detach_parse(pout, 0), _detach_place(qout, 0), None
        pb = self._pb
        ar = pb.pack(messages)
        pout, qout = eng.route(ar, pout=pb.pout, qout=pb.qout)
        fout = None
        if ((pout.status[:ar.n] & N.F_CHUNKED) != 0).any():
            fout = _detach_frame(eng.frame(ar, pout, reuse_arena=True, out=pb.fout, chunk_cap=pb.fout.chunk_cap), ar.n)
        return _detach_parse(pout, ar.n), _detach_place(qout, ar.n), fout

    def place_keys(self, keys):
        eng = self.eng
        ar = self._pb.pack(keys)
        out = eng.place(ar, out=self._pb.qout)
        return out.node[:len(keys)].copy()


class _Bag(object):
    pass


def _detach_parse(p, n):
    from .server.HttpParser import detach_parse
    return detach_parse(p, n)


def _detach_frame(f, n):
    from .server.HttpParser import detach_frame
    return detach_frame(f, n)


def _detach_place(q, n):
    b = _Bag()
    b.node, b.hash = q.node[:n].copy(), q.hash[:n].copy()
    return b


class ReactorShim(object):
    MAX_ROUNDS = 64          # pipelined requests peeled per connection per tick

    def __init__(self, router):
        self.router = router
        self.pending = {}                  # fd -> bytearray of bytes not yet consumed

    def feed(self, fd, data):
        """Called from the read handler instead of request.parse(data, len(data))."""
        if isinstance(data, str):
            data = data.encode('latin-1')
        self.pending.setdefault(fd, bytearray()).extend(data)

    def close(self, fd):
        self.pending.pop(fd, None)

    def tick(self):
        """One batch per round over every connection with pending bytes.  Returns a list of
        Routed in (round, fd) order.  Connections whose next request is incomplete keep their
        bytes; a request the reference would have raised on is returned with `.error` set and its
        connection's bytes are dropped."""
        done = []
        live = [fd for fd, buf in self.pending.items() if len(buf)]
        for _ in range(self.MAX_ROUNDS):
            if not live:
                break
            msgs = [bytes(self.pending[fd]) for fd in live]
            pout, qout, fout = self.router.route(msgs)
            late_keys, late_items, again = [], [], []
            # Per request only status, bytes consumed, URL span and node are needed: they are worked
            # out for the whole round with array arithmetic.  Whole, un-chunked requests (the bulk
            # of a tick) never get a Python parser object unless the caller asks for `.view`.
            n = len(live)
            st = np.asarray(pout.status[:n])
            code = st & 0xff
            plain = ((st & (N.F_MESSAGE_COMPLETE | N.F_IS_REQUEST | N.F_FIRST_LINE_DONE)) ==
                     (N.F_MESSAGE_COMPLETE | N.F_IS_REQUEST | N.F_FIRST_LINE_DONE)) & (code == N.S_OK) & \
                    ((st & (N.F_CHUNKED | N.F_CLEN_WIDE | N.F_KEY_TOO_LONG)) == 0)
            body = np.where((st & N.F_ON_BODY) == 0, np.asarray(pout.content_len[:n]), 0)
            used_v = (np.asarray(pout.first_line_len[:n]).astype(np.int64) + pout.headers_len[:n] + body).tolist()
            sp1_v, sp2_v = np.asarray(pout.sp1[:n]).tolist(), np.asarray(pout.sp2[:n]).tolist()
            node_v = np.asarray(qout.node[:n]).tolist()
            waiting = ((code == N.S_INCOMPLETE_LINE) | (code == N.S_INCOMPLETE_HEADERS) | (code == N.S_EMPTY) |
                       ((code == N.S_INCOMPLETE_BODY) & ((st & N.F_CHUNKED) == 0))).tolist()
            plain = plain.tolist()
            pending = self.pending
            for i, fd in enumerate(live):
                if plain[i]:
                    msg, used = msgs[i], used_v[i]
                    buf = pending[fd]
                    del buf[:used]
                    done.append(Routed(fd, None, msg[sp1_v[i] + 1:sp2_v[i]], node_v[i], None, msg[:used], (msg, pout, i, fout)))
                    if buf:
                        again.append(fd)
                    continue
                if waiting[i]:
                    continue                                   # wait for more bytes (next tick)
                v = MessageView(msgs[i], pout, i, fout)
                exc = v.exception()
                if exc is not None:
                    done.append(Routed(fd, v, None, None, exc(N.CODE_NAMES.get(v.code, '?')), msgs[i]))
                    self.pending[fd] = bytearray()
                    continue
                if not v.message_complete():
                    continue
                used = v.consumed()
                raw = msgs[i][:used]
                del self.pending[fd][:used]
                key = v.url()
                node = int(qout.node[i]) if key is not None else None
                item = Routed(fd, v, key, node, None, raw)
                if v.st & N.F_KEY_TOO_LONG:                    # pylibmc raises before hashing (Server.py:335)
                    item.node, item.error = None, ValueError('key too long, max is %d' % N.MAX_KEY_LEN)
                elif key is not None and node == 0xffffffff:   # completed by the chunk framer: place it now
                    late_keys.append(key)
                    late_items.append(item)
                done.append(item)
                if len(self.pending[fd]):
                    again.append(fd)
            if late_keys:
                for item, nd in zip(late_items, self.router.place_keys(late_keys)):
                    item.node = int(nd)
            live = again
        return done
