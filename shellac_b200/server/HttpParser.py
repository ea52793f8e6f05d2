"""HttpParser — the reference's class and method surface (HttpParser.py:43-351) over the device path.

    p = HttpParser()
    while not p.message_complete():
        c = p.parse(buf, len(buf))
        buf = buf[c:]
    p.headers(); p.body().read()

Same constructor, same `parse(data, length)` return values per call (`None` for length 0), same
getters, same exception TYPES at the same call.  The boundaries (first CRLF, CRLFCRLF, first-line
fields, header spans, chunk headers) come from the CUDA kernels through the C ABI; this class only
keeps the reference's per-connection buffering (HttpParser.py:153,174,209,247) and turns spans into
Python objects.  There is no CPU parser behind it: without the library / a GPU it raises.

`HttpParser.parse_many(messages)` is the batch entry: one device call for the whole list.

Type rule: feed `str` and the getters return `str` (latin-1), feed `bytes` and they return `bytes`
(a Python-2 `str` is bytes; the reference predates the split).
"""
import io
import zlib

import numpy as np

from .. import _native as N
from ..batch import MessageView

_FIRST_LINE_ERRORS = (N.E_FIRSTLINE_PARTS, N.E_VERSION_SPLIT, N.E_VERSION_FLOAT, N.E_STATUS_INT)


class DeviceBackend(object):
    """Runs single buffers through an Engine, reusing one pinned staging area."""

    def __init__(self, engine=None):
        if engine is None:
            from ..engine import Engine
            engine = Engine(device=0, max_msgs=1 << 16, max_bytes=1 << 26)
        self.eng = engine
        from ..engine import PinnedBatch
        self._pb = PinnedBatch(engine)
        self._ch = (engine.pinned(1, np.int32), engine.pinned(1, np.uint64), engine.pinned(1, np.uint8))

    def batch(self, bufs, frame=False):
        """One device call for a list of buffers (bytes): (parse rows, frame rows or None), detached
        from the pinned staging area, which is kept between calls and only grows."""
        eng, pb = self.eng, self._pb
        ar = pb.pack(bufs)
        po = eng.parse(ar, out=pb.pout)
        fo = None
        n = ar.n
        st = po.status[:n]
        chunked = frame and n and ((st & N.F_CHUNKED) != 0).any()
        # a message has body bytes when its headers are done and something follows them
        bodies = frame and n and (((st & N.F_HEADERS_COMPLETE) != 0) &
                                  (ar.len[:n] > po.first_line_len[:n] + po.headers_len[:n])).any()
        if chunked or bodies:
            fo = eng.frame(ar, po, reuse_arena=True, out=pb.fout, chunk_cap=pb.fout.chunk_cap)
        rows = detach_parse(po, n)
        if bodies:
            # N1: the bodies are gathered on the device (de-chunked / Content-Length slices, one
            # contiguous body per message) instead of being sliced out of the message on the host
            bo = eng.body(ar, out=pb.bout(ar.nbytes + 16 * n + 64))
            used = int(bo.body_used[0])
            rows.body_off, rows.body_len, rows.body = bo.body_off[:n].copy(), bo.body_len[:n].copy(), bo.body[:used].copy()
        return rows, (detach_frame(fo, n) if (fo is not None and chunked) else None)

    def views(self, bufs, frame=False):
        """MessageView per buffer (bytes)."""
        po, fo = self.batch(bufs, frame)
        return [MessageView(bufs[i], po, i, fo) for i in range(len(bufs))]

    def chunk_header(self, buf):
        hl, sz, ok = self.eng.chunk_header(self._pb.pack([buf]), out=self._ch)
        return int(hl[0]), int(sz[0]), bool(ok[0])


class _Rows(object):
    pass


def detach_parse(p, n):
    """Copies of the first n rows of a ParseOut and of the header records they point at.  (On the
    chunked large-batch path each chunk's records start at its own slice of the arena, so the
    records in use end at max(hdr_base + n_headers), not at the sum of the counts.)"""
    r = _Rows()
    for name, _ in N.PARSE_FIELDS:
        setattr(r, name, getattr(p, name)[:n].copy())
    end = int((r.hdr_base.astype(np.int64) + r.n_headers).max()) if n else 0
    r.hdr = p.hdr[:end].copy()
    return r


def detach_frame(f, n):
    r = _Rows()
    for name, _ in N.FRAME_FIELDS:
        setattr(r, name, getattr(f, name)[:n].copy())
    end = int((r.chunk_base.astype(np.int64) + r.n_chunks).max()) if n else 0
    r.chunk = f.chunk[:end].copy()
    return r


_default_backend = None


def default_backend():
    global _default_backend
    if _default_backend is None:
        _default_backend = DeviceBackend()
    return _default_backend


def set_default_backend(backend):
    """Tests on a machine without a GPU inject an oracle-backed backend here; the product never does."""
    global _default_backend
    _default_backend = backend


class HttpParser(object):

    def __init__(self, backend=None):
        self._backend = backend
        self._method = self._version = self._url = self._status = self._message = None
        self._headers = {}
        self._body = None
        self._buf = b''
        self._line = b''                  # the consumed first line: the device re-reads it with the headers
        self._is_request = True
        self._content_len = None
        self._chunked = False
        self._text = True
        self._inflate = zlib.decompressobj(31)            # HttpParser.py:58
        self._on_first_line, self._on_headers, self._on_body = True, False, False
        self._headers_complete = self._message_complete = False

    # ------------------------------------------------------------------ getters (HttpParser.py:67-109)
    def _be(self):
        return self._backend or default_backend()

    def _t(self, b):
        if b is None or not isinstance(b, (bytes, bytearray)):
            return b
        return bytes(b).decode('latin-1') if self._text else bytes(b)

    def method(self):
        return self._t(self._method)

    def url(self):
        return self._t(self._url)

    def status(self):
        return self._status

    def version(self):
        return self._version

    def message(self):
        return self._t(self._message)

    def headers(self):
        return self._headers

    def body(self):
        if self._body is None:
            self._body = io.StringIO() if self._text else io.BytesIO()
        return self._body

    def is_request(self):
        return self._is_request

    def is_response(self):
        return not self._is_request

    def headers_complete(self):
        return self._headers_complete

    def message_complete(self):
        return self._message_complete

    def _lit(self, s):
        return s if self._text else s.encode('latin-1')

    def keep_alive(self):
        return self._headers.get(self._lit('connection'), self._lit('close')).lower() == self._lit('keep-alive')

    def keep_alive_params(self):
        if not self.keep_alive():
            return (0, 1)
        ka = self._headers.get(self._lit('keep-alive'), self._lit('timeout=5, max=100'))
        kp = {}
        for item in ka.split(self._lit(', ')):
            k, v = item.split(self._lit('='))
            kp[k] = int(v)
        return (kp.get(self._lit('timeout'), 5), kp.get(self._lit('max'), 100))

    # ------------------------------------------------------------------ re-serialisation (HttpParser.py:111-138)
    def serialize(self):
        """The bytes `str(parser)` yields in the reference.  Header ORDER there is Python-2 dict
        hash order (HttpParser.py:23,136) and not reproducible; here it is arrival order."""
        L = lambda s: s.encode('latin-1') if isinstance(s, str) else s          # noqa: E731
        if self.is_request():
            s = b'%s %s HTTP/%s\r\n' % (L(self.method()), L(self.url()), b'%.1f' % self.version())
        else:
            s = b'HTTP/%s %d %s\r\n' % (b'%.1f' % self.version(), self.status(), L(self.message()))
        self._headers.pop(self._lit('transfer-encoding'), None)
        self._headers.pop(self._lit('content-length'), None)
        body = self.body()
        body.seek(0)
        b = L(body.read())
        if self._headers.get(self._lit('content-encoding'), self._lit('identity')) == self._lit('gzip'):
            zz = zlib.compressobj(6, zlib.DEFLATED, 31)
            b = zz.compress(b) + zz.flush()
        if len(b) != 0:
            self._headers[self._lit('content-length')] = len(b)
        h = b''
        for k, v in self._headers.items():
            name = b'-'.join(x.capitalize() for x in L(k).split(b'-'))
            if isinstance(v, list):
                val = b', '.join(L(x) for x in v)
            else:
                val = L(v) if isinstance(v, (str, bytes)) else b'%d' % v
            h += name + b': ' + val + b'\r\n'
        return s + h + b'\r\n' + b

    def __str__(self):
        return self.serialize().decode('latin-1')

    # ------------------------------------------------------------------ parse (HttpParser.py:140-232)
    def parse(self, data, length):
        """Parse data, return number of bytes consumed."""
        if length == 0:
            return None
        if isinstance(data, str):
            data = data.encode('latin-1')
        else:
            self._text = False
        if self._on_first_line:
            return self._parse_line(data, length)
        if self._on_headers:
            return self._parse_headers(data, length)
        if self._on_body:
            if self._chunked:
                return self._parse_chunked(data, length)
            return self._parse_sized(data, length)
        return 0

    def _parse_line(self, data, length):
        had = len(self._buf)
        self._buf += data
        v = self._be().views([self._buf])[0]
        self._seen = (self._buf, v)             # the device has tokenised these bytes: the header state may reuse it
        if v.code in (N.S_INCOMPLETE_LINE, N.S_EMPTY):
            return length                                                   # :156-157
        fl = int(v.p.first_line_len[0])
        self._buf = self._buf[:fl]                                          # :159
        self._is_request = v.is_request()
        self._version = v.version()
        if v.code in _FIRST_LINE_ERRORS:
            raise ValueError('malformed first line (%s)' % N.CODE_NAMES[v.code])   # :307-319
        self._method, self._url = v.method(), v.url()
        self._status, self._message = v.status(), v.message()
        self._line, self._buf = self._buf, b''
        self._on_first_line, self._on_headers = False, True
        return fl - had                                                     # :163

    def _parse_headers(self, data, length):
        had = len(self._buf)
        self._buf += data
        full = self._line + self._buf
        seen = getattr(self, '_seen', None)
        if seen is not None and seen[0] == full:        # the caller handed back the rest of what the first-line call saw:
            v = seen[1]                                 # one device round trip per message instead of two
        else:
            v = self._be().views([full])[0]
        self._seen = None
        if v.code == N.S_INCOMPLETE_HEADERS:
            return length                                                   # :177-178
        if v.code == N.E_LIB_HDR_TOO_LARGE:
            raise OverflowError('header block beyond 64 KiB (library limit)')
        hl = int(v.p.headers_len[0])
        self._buf = self._buf[:hl]                                          # :180
        v.text = self._text
        for k, val in v.header_items():                                     # :327-337
            if k in self._headers:
                if isinstance(self._headers[k], list):
                    self._headers[k].append(val)
                else:
                    self._headers[k] = [self._headers[k], val]
            else:
                self._headers[k] = val
        if v.code in (N.E_HDR_NO_SEP, N.E_HDR_MULTI_SEP):
            raise ValueError('malformed header line (%s)' % N.CODE_NAMES[v.code])  # :328
        self._buf = b''
        self._headers_complete, self._on_headers, self._on_body = True, False, True
        if self._method in (b'GET', b'HEAD'):                               # :191
            self._message_complete = True
        self._chunked = v.chunked()                                         # :194
        if not self._chunked:
            if v.code == N.E_CLEN_INT:
                raise ValueError('invalid literal for int() in content-length')    # :197
            if v.code == N.E_CLEN_LIST:
                raise TypeError('int() argument must be a string, not list')       # :197
            self._content_len = v.content_len()
            if self._content_len == 0:                                      # :198-200
                self._on_body = False
                self._message_complete = True
        return hl - had                                                     # :184

    def _parse_sized(self, data, length):                                   # :206-223
        if length < self._content_len:
            self._buf += data
            self._content_len -= length
            return length
        self._buf += data[:self._content_len]
        self._write_body()
        self._flush_body()
        self._buf = b''
        used = self._content_len
        self._content_len = 0
        self._on_body = False
        self._message_complete = True
        return used

    def _parse_chunked(self, data, length):                                 # :243-301
        had = len(self._buf)
        self._buf += data
        used = 0
        if self._content_len is None:
            idx, size, ok = self._be().chunk_header(self._buf)              # CHUNK_HEADER_RX.match, :253
            if idx < 0:
                return length
            if not ok:
                raise ValueError('invalid literal for int() with base 16')  # :239-241
            if size >= (1 << 63):        # the device saturates; Python carries the exact long (:239-241)
                size = int(self._buf[:idx].strip().split(b';')[0], 16)
            self._content_len = size
            used = idx - had
            self._buf = self._buf[idx:]
            left = len(self._buf)
            if self._content_len == 0:                                      # :268-276
                if self._buf.startswith(b'\r\n'):
                    used += 2
                    self._flush_body()
                    self._on_body = False
                    self._message_complete = True
                self._buf = b''
                return used
        elif self._content_len == 0:                                        # :278-287
            if self._buf.startswith(b'\r\n'):
                used = 2 - had
                self._flush_body()
                self._buf = b''
                self._on_body = False
                self._message_complete = True
                return used
            return length
        else:
            left = length
        if left < self._content_len:                                        # :291-294
            self._content_len -= left
            return used + left
        self._buf = self._buf[:had + self._content_len]                     # :296 (stale `had` quirk kept)
        self._write_body()
        self._buf = b''
        used += self._content_len
        self._content_len = None
        return used

    # ------------------------------------------------------------------ body sink (HttpParser.py:340-351)
    def _gzip(self):
        return self._headers.get(self._lit('content-encoding'), self._lit('identity')) == self._lit('gzip')

    def _emit(self, raw):
        body = self.body()
        pos = body.tell()
        body.seek(0, 2)
        body.write(raw.decode('latin-1') if self._text else raw)
        body.seek(pos)

    def _write_body(self):
        self._emit(self._inflate.decompress(self._buf) if self._gzip() else self._buf)

    def _flush_body(self):
        if self._gzip():
            self._emit(self._inflate.flush())

    # ------------------------------------------------------------------ batch entry
    @classmethod
    def parse_many(cls, messages, backend=None):
        """One device call for a list of whole messages; returns a parser object per message in
        the state the reference object is in after being driven over that message.  A message
        whose parse raised in the reference carries the exception in `.error` instead.  The
        objects are cheap: each turns its row of the batch result into Python values on first use."""
        be = backend or default_backend()
        text = [isinstance(m, str) for m in messages]
        any_text = any(text)
        bufs = [m.encode('latin-1') if isinstance(m, str) else bytes(m) for m in messages] if any_text else \
               [m if type(m) is bytes else bytes(m) for m in messages]
        po, fo = be.batch(bufs, frame=True) if hasattr(be, 'batch') else _rows_of_views(be, bufs)
        mk = _LazyParser
        if not any_text:
            return [mk(be, po, fo, bufs[i], i, False) for i in range(len(bufs))]
        return [mk(be, po, fo, bufs[i], i, text[i]) for i in range(len(bufs))]


def _rows_of_views(be, bufs):
    """Backends that only offer views() (the oracle-backed one of the CPU tests)."""
    vs = be.views(bufs, frame=True)
    return (vs[0].p, vs[0].f) if vs else (None, None)


class _LazyParser(HttpParser):
    """A parser object of parse_many: row `i` of a batch result.  Nothing is converted until an
    attribute is asked for; then the whole object is filled in once (so every getter of the
    reference's surface, and the incremental state flags, behave as on a driven parser)."""

    def __init__(self, backend, po, fo, buf, i, text):   # noqa: super().__init__ is what _fill replaces
        d = self.__dict__
        d['_backend'], d['_lazy'] = backend, (po, fo, buf, i, text)

    def __getattr__(self, name):
        lazy = self.__dict__.get('_lazy')
        if lazy is None or name.startswith('__'):
            raise AttributeError(name)
        self._fill(lazy)
        return object.__getattribute__(self, name)

    def _fill(self, lazy):
        po, fo, buf, i, is_text = lazy
        self.__dict__['_lazy'] = None
        HttpParser.__init__(self, backend=self._backend)
        v = MessageView(buf, po, i, fo, text=is_text)
        self._text = is_text
        self._is_request, self._version = v.is_request(), v.version()
        un = (lambda x: x.encode('latin-1') if isinstance(x, str) else x)
        self._method, self._url, self._message, self._status = un(v.method()), un(v.url()), un(v.message()), v.status()
        self._headers = v.headers()
        self._headers_complete, self._message_complete = v.headers_complete(), v.message_complete()
        self._chunked, self._on_body = v.chunked(), v.on_body()
        self._on_first_line = not (v.st & N.F_FIRST_LINE_DONE)
        self._on_headers = bool(v.st & N.F_FIRST_LINE_DONE) and not self._headers_complete
        self._body = v.body()
        self.calls = v.calls()
        exc = v.exception()
        self.error = exc() if exc else None
