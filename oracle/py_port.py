"""py_port.py — CPython restatement of the reference's hot path on `bytes`.   TEST INFRASTRUCTURE.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this; the shellac_b200 package never does.

Why it exists next to the C oracle: the reference IS CPython (HttpParser.py is pure Python), so the
honest CPU baseline for "the reference's own CPU implementation of the path" is the same algorithm
in the same interpreter: the same sequence of `+=`, `find`, `split`, `strip`, `lower` and dict
operations per message as /root/reference/src/python/shellac/server/HttpParser.py:140-337, the
key = url() of Server.py:327 and, for placement (pylibmc/libmemcached are third-party C and absent),
hashlib.md5 + bisect as SURVEY.md section 8c restates libmemcached 0.44.  /root/reference does not
exist on the GPU box, so the reference file itself cannot be timed there; this port is pinned
against the reference's golden vectors in tests/test_oracle_golden.py::test_py_port_*.

Operating on bytes gives Python-2 `str` semantics for free: bytes.strip()/lower()/split() are
ASCII-only exactly like py2 str in the C locale.  int()/float() of bytes accept `_` separators in
Python 3 (py2 does not): fenced out of every corpus.
"""
import bisect
import hashlib
import io
import re
import struct
import zlib

# HttpParser.py:39-40
_CHUNK_RX = re.compile(br'(\r\n)?[a-z0-9]+(;[a-z0-9]+="?[a-z0-9\-_]+"?)?\r\n', re.IGNORECASE)

ST_LINE, ST_HEADERS, ST_BODY, ST_DONE = 0, 1, 2, 3


class PortParser(object):
    """Same observable behaviour as the reference HttpParser, `bytes` in / `bytes` out."""

    __slots__ = ('method', 'version', 'url', 'status', 'reason', 'headers', 'body', 'pending', 'is_request',
                 'content_len', 'chunked', 'state', 'headers_complete', 'message_complete', 'inflater')

    def __init__(self):
        self.method = self.version = self.url = self.status = self.reason = None
        self.headers = {}
        self.body = io.BytesIO()
        self.pending = b''
        self.is_request = True
        self.content_len = None
        self.chunked = False
        self.state = ST_LINE
        self.headers_complete = self.message_complete = False
        self.inflater = zlib.decompressobj(31)                    # :58

    # -- parse() :140-232 ------------------------------------------------------------------
    def parse(self, data, length):
        if length == 0:
            return None
        st = self.state
        if st == ST_LINE:
            had = len(self.pending)
            self.pending += data
            at = self.pending.find(b'\r\n')
            if at < 0:
                return length
            self.pending = self.pending[:at + 2]
            self._first_line()
            used = len(self.pending) - had
            self.pending = b''
            self.state = ST_HEADERS
            return used
        if st == ST_HEADERS:
            had = len(self.pending)
            self.pending += data
            at = self.pending.find(b'\r\n\r\n')
            if at < 0:
                return length
            self.pending = self.pending[:at + 4]
            self._header_block()
            used = len(self.pending) - had
            self.pending = b''
            self.headers_complete = True
            self.state = ST_BODY
            if self.method in (b'GET', b'HEAD'):
                self.message_complete = True
            self.chunked = self.headers.get(b'transfer-encoding', b'none') == b'chunked'
            if not self.chunked:
                self.content_len = int(self.headers.get(b'content-length', 0))
                if self.content_len == 0:
                    self.state = ST_DONE
                    self.message_complete = True
            return used
        if st == ST_BODY:
            if self.chunked:
                return self._chunked(data, length)
            if length < self.content_len:
                self.pending += data
                self.content_len -= length
                return length
            self.pending += data[:self.content_len]
            self._emit_body()
            self._flush()
            self.pending = b''
            used = self.content_len
            self.content_len = 0
            self.state = ST_DONE
            self.message_complete = True
            return used
        return 0

    # -- :304-321 ----------------------------------------------------------------------------
    def _first_line(self):
        a, b, c = self.pending.rstrip().split(b' ', 2)
        if a.startswith(b'HT'):
            self.is_request = False
            _, v = a.split(b'/')
            self.version = float(v)
            self.status = int(b)
            self.reason = c
        else:
            self.is_request = True
            _, v = c.split(b'/')
            self.version = float(v)
            self.method = a
            self.url = b

    # -- :323-337 ----------------------------------------------------------------------------
    def _header_block(self):
        hs = self.headers
        for line in self.pending.strip().split(b'\r\n'):
            key, value = line.split(b': ')
            value = value.strip()
            key = key.lower()
            if key in hs:
                if isinstance(hs[key], list):
                    hs[key].append(value)
                else:
                    hs[key] = [hs[key], value]
            else:
                hs[key] = value

    # -- :234-301 ----------------------------------------------------------------------------
    def _chunked(self, data, length):
        had = len(self.pending)
        self.pending += data
        used = 0
        if self.content_len is None:
            m = _CHUNK_RX.match(self.pending)
            if m is None:
                return length
            at = len(m.group())
            size = self.pending[:at].strip()
            self.content_len = int(size.split(b';')[0] if b';' in size else size, 16)
            used = at - had
            self.pending = self.pending[at:]
            left = len(self.pending)
            if self.content_len == 0:
                if self.pending.startswith(b'\r\n'):
                    used += 2
                    self._flush()
                    self.state = ST_DONE
                    self.message_complete = True
                self.pending = b''
                return used
        elif self.content_len == 0:
            if self.pending.startswith(b'\r\n'):
                used = 2 - had
                self._flush()
                self.pending = b''
                self.state = ST_DONE
                self.message_complete = True
                return used
            return length
        else:
            left = length
        if left < self.content_len:
            self.content_len -= left
            return used + left
        self.pending = self.pending[:had + self.content_len]
        self._emit_body()
        self.pending = b''
        used += self.content_len
        self.content_len = None
        return used

    # -- :340-351 ----------------------------------------------------------------------------
    def _gzip(self):
        return self.headers.get(b'content-encoding', b'identity') == b'gzip'

    def _emit_body(self):
        pos = self.body.tell()
        self.body.seek(0, 2)
        self.body.write(self.inflater.decompress(self.pending) if self._gzip() else self.pending)
        self.body.seek(pos)

    def _flush(self):
        if self._gzip():
            self.body.write(self.inflater.flush())


def drive(msg):
    """test/HttpParserTests.py:36-39, stopping when the data runs out."""
    p = PortParser()
    data, calls, exc = msg, [], None
    try:
        while data and not p.message_complete:
            c = p.parse(data, len(data))
            calls.append(c)
            data = data[c:]
    except Exception as e:          # noqa: BLE001 - the type is the observation
        exc = type(e).__name__
    return p, calls, exc


# ------------------------------------------------------------------------------- placement
def build_ring(hosts, ports=None, weights=None, mode=0):
    """libmemcached 0.44 update_continuum, restated (see oracle/shl_oracle.c)."""
    import numpy as np
    ports = ports or [11211] * len(hosts)
    weights = weights or [1] * len(hosts)
    pts = []
    W = np.float32(sum(weights))
    for s, (h, p, w) in enumerate(zip(hosts, ports, weights)):
        reps, per = 100, 1
        if mode == 1:
            t = np.float32(np.float32(w) / W) * np.float32(160)
            t = np.float32(np.float32(t / np.float32(4)) * np.float32(len(hosts)))
            reps, per = int(np.floor(np.float32(np.float64(t) + 0.0000000001))), 4
        for i in range(reps):
            label = ('%s-%d' % (h, i)) if p == 11211 else ('%s:%d-%d' % (h, p, i))
            d = hashlib.md5(label.encode()).digest()
            for a in range(per):
                pts.append((struct.unpack_from('<I', d, 4 * a)[0], s))
    pts.sort(key=lambda x: x[0])
    return [v for v, _ in pts], [n for _, n in pts]


def place(values, nodes, n_hosts, key):
    h = struct.unpack_from('<I', hashlib.md5(key).digest())[0]
    j = bisect.bisect_left(values, h)
    return 0 if n_hosts == 1 else nodes[0 if j == len(values) else j]


def route_messages(msgs, values, nodes, n_hosts):
    """The timed unit of the CPU baseline: parse each message to completion, key = url(), place."""
    hist = [0] * n_hosts
    done = 0
    for m in msgs:
        p = PortParser()
        data = m
        while data and not p.message_complete:
            c = p.parse(data, len(data))
            data = data[c:]
        if p.message_complete and p.url is not None:
            hist[place(values, nodes, n_hosts, p.url)] += 1
            done += 1
    return done, hist


# ---------------------------------------------------------------------------------------------
# Memcached binary protocol, request side (what pylibmc.Client(..., binary=True) sends for
# get(key) / set(key, value, time=ttl), Server.py:81-83, 335, 432).  pylibmc / libmemcached are
# absent; restated from the published protocol: 24-byte header, network byte order.
def mc_get_frame(key, opaque):
    return struct.pack('>BBHBBHIIQ', 0x80, 0x00, len(key), 0, 0, 0, len(key), opaque, 0) + key


def mc_set_header(key, opaque, value_len, ttl, flags=0):
    """Header + extras + key of a Set; the value's bytes follow on the wire."""
    return struct.pack('>BBHBBHIIQ', 0x80, 0x01, len(key), 8, 0, 0, 8 + len(key) + value_len, opaque, 0) + \
        struct.pack('>II', flags, ttl) + key


def serialize(p):
    """HttpParser.__str__ (:111-138) on a driven PortParser, as bytes.  Header order is arrival
    order here (Python-2 dict hash order in the reference): compare the lines as a multiset."""
    if p.is_request:
        s = b'%s %s HTTP/%s\r\n' % (p.method, p.url, b'%.1f' % p.version)
    else:
        s = b'HTTP/%s %d %s\r\n' % (b'%.1f' % p.version, p.status, p.reason)
    p.headers.pop(b'transfer-encoding', None)
    p.headers.pop(b'content-length', None)
    p.body.seek(0)
    b = p.body.read()
    if p.headers.get(b'content-encoding', b'identity') == b'gzip':
        zz = zlib.compressobj(6, zlib.DEFLATED, 31)
        b = zz.compress(b) + zz.flush()
    if len(b) != 0:
        p.headers[b'content-length'] = len(b)
    h = b''.join(b'%s: %s\r\n' % (b'-'.join(x.capitalize() for x in k.split(b'-')),
                                  b', '.join(v) if isinstance(v, list) else (v if isinstance(v, bytes) else b'%d' % v))
                 for k, v in p.headers.items())
    return s + h + b'\r\n' + b
