"""Host-side views of the struct-of-arrays results the device writes.

`ParseOut` / `FrameOut` / `PlaceOut` own the numpy arrays a *_batch call fills (pinned when they
come from an Engine) and hand the C ABI a struct of their addresses.  `MessageView` turns one row
back into what the reference's HttpParser object would hold after being driven over that message
(HttpParser.py:67-109 accessors), so tests can compare row by row with the reference.
"""
import io
import zlib

import numpy as np

from . import _native as N


def _np_alloc(n, dtype):
    return np.zeros(max(int(n), 1), dtype=dtype)


class Arena(object):
    """Offset+length indexed byte arena: message i is data[off[i] : off[i]+len[i]].

    This is the batch-side replacement of the strings the reference passes around
    (`data = conn.recv(4096)`, Server.py:311,394): one contiguous buffer per batch, so a single
    H2D copy moves every message.
    """

    def __init__(self, data, off, length):
        self.data, self.off, self.len = data, off, length
        self.n = int(len(off))
        self._nbytes = None

    @classmethod
    def pack(cls, messages, alloc=_np_alloc, align=1):
        """Pack an iterable of bytes-like messages back to back (each start aligned to `align`)."""
        msgs = [m.encode('latin-1') if isinstance(m, str) else bytes(m) for m in messages]
        n = len(msgs)
        lens = np.fromiter((len(m) for m in msgs), dtype=np.uint32, count=n)
        step = (lens.astype(np.uint64) + np.uint64(align - 1)) // np.uint64(align) * np.uint64(align)
        off = np.zeros(n, dtype=np.uint64)
        if n > 1:
            np.cumsum(step[:-1], out=off[1:])
        total = int(off[-1] + lens[-1]) if n else 0
        data = alloc(total + 64, np.uint8)          # +64: the device reads whole 16-byte vectors
        if align == 1:
            blob = b''.join(msgs)
            data[:total] = np.frombuffer(blob, dtype=np.uint8)
        else:
            for m, o in zip(msgs, off):
                data[int(o):int(o) + len(m)] = np.frombuffer(m, dtype=np.uint8)
        a_off, a_len = alloc(n, np.uint64), alloc(n, np.uint32)
        a_off[:n], a_len[:n] = off, lens
        return cls(data, a_off[:n] if n else a_off[:0], a_len[:n] if n else a_len[:0])

    @property
    def nbytes(self):
        """One past the last message byte (messages need not be stored in index order)."""
        if not self.n:
            return 0
        if self._nbytes is None:
            self._nbytes = int((self.off.astype(np.uint64) + self.len.astype(np.uint64)).max())
        return self._nbytes

    def message(self, i):
        o = int(self.off[i])
        return self.data[o:o + int(self.len[i])].tobytes()


class ParseOut(object):
    def __init__(self, n, hdr_cap, alloc=_np_alloc, lkeys_cap=0):
        """lkeys_cap > 0 also asks for the header names lower-cased on the device (key.lower(),
        HttpParser.py:330): `lkeys` arena + `lkey_base[i]`, see include/shellac_b200.h."""
        self.n, self.hdr_cap, self.lkeys_cap = int(n), int(hdr_cap), int(lkeys_cap)
        for name, dt in N.PARSE_FIELDS:
            setattr(self, name, alloc(n, dt))
        self.hdr = alloc(hdr_cap, N.HDR_DTYPE)
        self.hdr_used = alloc(1, np.uint64)
        self.lkeys = self.lkeys_used = self.lkey_base = None
        if lkeys_cap:
            self.lkeys, self.lkeys_used, self.lkey_base = alloc(lkeys_cap, np.uint8), alloc(1, np.uint64), alloc(n, np.uint32)

    def lower_keys(self, i):
        """Message i's header names as the device lower-cased them, in record order."""
        b, k = int(self.hdr_base[i]), int(self.n_headers[i])
        p, out = int(self.lkey_base[i]), []
        for kl in self.hdr['key_len'][b:b + k].tolist():
            out.append(self.lkeys[p:p + kl].tobytes())
            p += kl
        return out

    def struct(self):
        """The C struct of this object's array addresses (built once: the arrays never move)."""
        s = getattr(self, '_struct', None)
        if s is None:
            s = N.ParseOutStruct()
            for name, _ in N.PARSE_FIELDS:
                setattr(s, name, getattr(self, name).ctypes.data)
            s.hdr, s.hdr_cap, s.hdr_used = self.hdr.ctypes.data, self.hdr_cap, self.hdr_used.ctypes.data
            if self.lkeys is not None:
                s.lkeys, s.lkeys_cap = self.lkeys.ctypes.data, self.lkeys_cap
                s.lkeys_used, s.lkey_base = self.lkeys_used.ctypes.data, self.lkey_base.ctypes.data
            self._struct = s
        return s

    def code(self):
        return self.status[:self.n] & 0xff


class FrameOut(object):
    def __init__(self, n, chunk_cap, alloc=_np_alloc):
        self.n, self.chunk_cap = int(n), int(chunk_cap)
        for name, dt in N.FRAME_FIELDS:
            setattr(self, name, alloc(n, dt))
        self.chunk = alloc(chunk_cap, N.CHUNK_DTYPE)
        self.chunk_used = alloc(1, np.uint64)

    def struct(self):
        s = N.FrameOutStruct()
        for name, _ in N.FRAME_FIELDS:
            setattr(s, name, getattr(self, name).ctypes.data)
        s.chunk, s.chunk_cap, s.chunk_used = self.chunk.ctypes.data, self.chunk_cap, self.chunk_used.ctypes.data
        return s


class BodyOut(object):
    def __init__(self, n, body_cap, alloc=_np_alloc):
        self.n, self.body_cap = int(n), int(body_cap)
        self.body_off, self.body_len = alloc(n, np.uint64), alloc(n, np.uint64)
        self.body = alloc(body_cap, np.uint8)
        self.body_used = alloc(1, np.uint64)

    def struct(self):
        s = N.BodyOutStruct()
        s.body_off, s.body_len = self.body_off.ctypes.data, self.body_len.ctypes.data
        s.body, s.body_cap, s.body_used = self.body.ctypes.data, self.body_cap, self.body_used.ctypes.data
        return s

    def get(self, i):
        o = int(self.body_off[i])
        return self.body[o:o + int(self.body_len[i])].tobytes()


class PlaceOut(object):
    def __init__(self, n, n_nodes=0, alloc=_np_alloc, want_md5=True):
        self.n = int(n)
        self.md5 = alloc(16 * max(n, 1), np.uint8) if want_md5 else None
        self.hash = alloc(n, np.uint32)
        self.node = alloc(n, np.uint32)
        self.hist = alloc(n_nodes, np.uint64) if n_nodes else None

    def struct(self):
        s = getattr(self, '_struct', None)
        if s is None:
            s = N.PlaceOutStruct()
            s.md5 = self.md5.ctypes.data if self.md5 is not None else None
            s.hash, s.node = self.hash.ctypes.data, self.node.ctypes.data
            s.hist = self.hist.ctypes.data if self.hist is not None else None
            self._struct = s
        return s

    def digest(self, i):
        return self.md5[16 * i:16 * i + 16].tobytes()


_GZIP_WBITS = 31        # zlib.decompressobj(31), HttpParser.py:58


class MessageView(object):
    """What the reference parser holds after `while data and not complete: parse(...)` over
    message i of a batch.  `text=True` returns latin-1 `str` (one char per byte) instead of bytes,
    which is how the Python-3-run reference oracle sees the same data."""

    def __init__(self, msg, pout, i, fout=None, text=False):
        self.m, self.p, self.i, self.f, self.text = bytes(msg), pout, int(i), fout, text
        self.st = int(pout.status[i])

    # -- helpers
    def _s(self, a, b):
        x = self.m[a:b]
        return x.decode('latin-1') if self.text else x

    def _lit(self, s):
        return s if self.text else s.encode('latin-1')

    @property
    def code(self):
        return self.st & 0xff

    def exception(self):
        """The exception type the reference raises while parsing this message, or None."""
        c = self.code
        if self.f is not None and self.st & N.F_CHUNKED and int(self.f.fstatus[self.i]) == N.E_CHUNK_HEX:
            return ValueError
        return N.CODE_EXC.get(c)

    # -- HttpParser.py:67-98
    def is_request(self):
        return bool(self.st & N.F_IS_REQUEST)

    def is_response(self):
        return not self.is_request()

    def method(self):
        if self.st & N.F_FIRST_LINE_DONE and self.is_request():
            return self._s(0, int(self.p.sp1[self.i]))
        return None

    def url(self):
        if self.st & N.F_FIRST_LINE_DONE and self.is_request():
            return self._s(int(self.p.sp1[self.i]) + 1, int(self.p.sp2[self.i]))
        return None

    def version_text(self):
        vc = (self.st >> N.F_VERSION_SHIFT) & 3
        if vc == 3:
            return None
        a = int(self.p.slash[self.i]) + 1
        b = int(self.p.fl_end[self.i]) if self.is_request() else int(self.p.sp1[self.i])
        return self.m[a:b]

    def version(self):
        vc = (self.st >> N.F_VERSION_SHIFT) & 3
        if vc == 3:
            return None
        if vc == 0:
            return 1.0
        if vc == 1:
            return 1.1
        return float(self.version_text().decode('latin-1'))   # device validated the literal

    def status(self):
        if not (self.st & N.F_FIRST_LINE_DONE) or self.is_request():
            return None
        if self.st & N.F_STATUS_WIDE:
            return int(self.m[int(self.p.sp1[self.i]) + 1:int(self.p.sp2[self.i])].decode('latin-1'))
        return int(self.p.status_code[self.i])

    def message(self):
        if self.st & N.F_FIRST_LINE_DONE and self.is_response():
            return self._s(int(self.p.sp2[self.i]) + 1, int(self.p.fl_end[self.i]))
        return None

    def header_items(self):
        """(lower(key), stripped value) in arrival order (HttpParser.py:327-330)."""
        b, n = int(self.p.hdr_base[self.i]), int(self.p.n_headers[self.i])
        m = self.m
        recs = np.ascontiguousarray(self.p.hdr[b:b + n]).view(np.uint16).reshape(-1, 4).tolist()     # one conversion per message
        if not self.text:                 # bytes.lower(): ASCII only, like py2 str.lower()
            return [(m[ko:ko + kl].lower(), m[vo:vo + vl]) for ko, kl, vo, vl in recs]
        return [(m[ko:ko + kl].lower().decode('latin-1'), m[vo:vo + vl].decode('latin-1')) for ko, kl, vo, vl in recs]

    def _lower(self, k):
        k = k.lower()                     # bytes.lower(): ASCII only, like py2 str.lower()
        return k.decode('latin-1') if self.text else k

    def headers(self):
        """dict with duplicate keys folded into lists in arrival order (HttpParser.py:331-337)."""
        d = getattr(self, '_hd', None)
        if d is not None and self._hd_text == self.text:
            return d
        d = {}
        for k, v in self.header_items():
            if k in d:
                if isinstance(d[k], list):
                    d[k].append(v)
                else:
                    d[k] = [d[k], v]
            else:
                d[k] = v
        self._hd, self._hd_text = d, self.text
        return d

    def headers_complete(self):
        return bool(self.st & N.F_HEADERS_COMPLETE)

    def chunked(self):
        return bool(self.st & N.F_CHUNKED)

    def _frame_done(self):
        return (self.f is not None and self.chunked() and int(self.f.fstatus[self.i]) == N.S_OK)

    def message_complete(self):
        return bool(self.st & N.F_MESSAGE_COMPLETE) or self._frame_done()

    def on_body(self):
        if self._frame_done():
            return False
        return bool(self.st & N.F_ON_BODY)

    def body_off(self):
        return int(self.p.first_line_len[self.i]) + int(self.p.headers_len[self.i])

    def content_len(self):
        """int(headers.get('content-length', 0)) as the header state left it (HttpParser.py:197)."""
        if self.st & N.F_CLEN_WIDE:
            for k, v in self.header_items():
                if k == self._lit('content-length'):
                    return int(v if self.text else v.decode('latin-1'))
        return int(self.p.content_len[self.i])

    def _body_call(self):
        """True when the driver loop reaches the body state (a third parse() call exists)."""
        if not self.headers_complete() or self.code >= 16:
            return False
        if self.is_request() and self.m[:int(self.p.sp1[self.i])] in (b'GET', b'HEAD'):
            return False
        if not self.chunked() and self.content_len() == 0:
            return False
        return len(self.m) > self.body_off()

    def chunks(self):
        if self.f is None or not self.chunked() or int(self.f.fstatus[self.i]) == N.FS_NOT_CHUNKED:
            return []
        b, n = int(self.f.chunk_base[self.i]), int(self.f.n_chunks[self.i])
        return [(int(c['hdr_off']), int(c['hdr_len']), int(c['data_off']), int(c['data_len']))
                for c in self.f.chunk[b:b + n]]

    def raw_body(self):
        """Body bytes before content-decoding, as the reference accumulates them
        (HttpParser.py:213-214 and :296-297)."""
        if not self._body_call():
            return b''
        rows = self.p
        if getattr(rows, 'body', None) is not None:             # gathered on the device (shl_body_batch)
            o = int(rows.body_off[self.i])
            return rows.body[o:o + int(rows.body_len[self.i])].tobytes()
        if self.chunked():
            return b''.join(self.m[d:d + n] for (_, _, d, n) in self.chunks())
        if not self.message_complete():
            return b''                                  # partial data stays in _buf (:209)
        return self.m[self.body_off():][:self.content_len()]     # python slice semantics (:213)

    def body(self):
        raw = self.raw_body()
        enc = self.headers().get(self._lit('content-encoding'), self._lit('identity'))
        if enc == self._lit('gzip') and raw:            # HttpParser.py:343-344, host-side inflate
            z = zlib.decompressobj(_GZIP_WBITS)
            raw = z.decompress(raw) + (z.flush() if self.message_complete() else b'')
        return io.StringIO(raw.decode('latin-1')) if self.text else io.BytesIO(raw)

    def calls(self):
        """The nb_parsed sequence the reference's parse() returns over this message."""
        c, L = self.code, len(self.m)
        if c == N.S_EMPTY:
            return []
        if c == N.S_INCOMPLETE_LINE:
            return [L]
        if c in (N.E_FIRSTLINE_PARTS, N.E_VERSION_SPLIT, N.E_VERSION_FLOAT, N.E_STATUS_INT):
            return []
        fl = int(self.p.first_line_len[self.i])
        out = [fl]
        if c == N.S_INCOMPLETE_HEADERS:
            if L > fl:
                out.append(L - fl)
            return out
        if not self.headers_complete() or c in (N.E_CLEN_INT, N.E_CLEN_LIST):
            return out
        out.append(int(self.p.headers_len[self.i]))
        if not self._body_call():
            return out
        rem = L - self.body_off()
        if not self.chunked():
            cl = self.content_len()
            out.append(rem if rem < cl else cl)
            return out
        q = self.body_off()
        for (ho, hl, do, dl) in self.chunks():
            out.append(hl + dl)
            q = do + dl
        if int(self.f.fstatus[self.i]) == N.E_CHUNK_HEX:
            return out
        tail, last = int(self.f.body_end[self.i]) - q, int(self.f.last_call[self.i])
        if tail > 0:
            out += [tail] if last == tail else [tail - last, last]
        return out

    def consumed(self):
        return sum(self.calls())

    def keep_alive(self):                                 # HttpParser.py:100-101
        v = self.headers().get(self._lit('connection'), self._lit('close'))
        return v.lower() == self._lit('keep-alive')

    def keep_alive_params(self):                          # HttpParser.py:103-109
        if not self.keep_alive():
            return (0, 1)
        ka = self.headers().get(self._lit('keep-alive'), self._lit('timeout=5, max=100'))
        kp = {}
        for item in ka.split(self._lit(', ')):
            k, v = item.split(self._lit('='))
            kp[k] = int(v)
        return (kp.get(self._lit('timeout'), 5), kp.get(self._lit('max'), 100))
