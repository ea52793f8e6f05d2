"""StreamBuf — the reference's ack-pointer byte queue (StreamBuf.py:35-78), same ten methods.

It is what carries bytes into and out of the hot path (Server.py:339,361,369,420).  Here the
queue is a `bytearray` plus a cursor, so `write` is amortised O(1) and a batch packer can take a
zero-copy `memoryview` of the unsent tail (`window()`) instead of the fresh string copy the
reference's `read()` makes every time (StreamBuf.py:56-57).  Like the reference it has "no
seatbelt": `ack`/`seek` are not bounds-checked (StreamBuf.py:27).  Not thread safe (:28).

The type that comes out follows the type that went in: feed `str` and `read()`/`buffer()` return
`str` (latin-1, one char per byte, which is what a Python-2 `str` is); feed `bytes`, get `bytes`.
"""


class StreamBuf(object):

    def __init__(self, data=None):
        self._store = bytearray()
        self._cursor = 0
        self._closed = False
        self._written = False
        self._text = True            # until told otherwise: '' == '' holds for the empty buffer
        if data:
            self.write(data)

    def _out(self, b):
        return bytes(b).decode('latin-1') if self._text else bytes(b)

    def write(self, data):                       # StreamBuf.py:46-48
        self._written = True
        if isinstance(data, str):
            data = data.encode('latin-1')
        else:
            self._text = False
        self._store += data

    def ack(self, bytes):                        # :50-51 (argument name kept)
        self._cursor += bytes

    def seek(self, pos):                         # :53-54
        self._cursor = pos

    def read(self):                              # :56-57
        return self._out(self._store[self._cursor:])

    def window(self):
        """Zero-copy view of the unsent tail, for packing into an arena."""
        return memoryview(self._store)[self._cursor:]

    def close(self):                             # :59-60
        self._closed = True

    def buffer(self):                            # :62-63
        return self._out(self._store)

    def clear(self):                             # :65-69
        self._store = bytearray()
        self._cursor = 0
        self._closed = False
        self._written = False

    def complete(self):                          # :71-72
        return self.closed() and self._cursor >= len(self._store)

    def closed(self):                            # :74-75
        return self._closed

    def ready(self):                             # :77-78
        return self._written
