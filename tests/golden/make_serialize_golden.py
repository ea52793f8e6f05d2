#!/usr/bin/env python
"""Golden vectors for N2, message re-serialisation: str(parser) of the UNMODIFIED reference
(HttpParser.__str__, /root/reference/src/python/shellac/server/HttpParser.py:111-138) over every
message of parser_golden.json.gz that the reference parses to completion without an exception.

TEST INFRASTRUCTURE; runs only where /root/reference is mounted.  Under CPython 3 `__str__` stops at
`dict.viewitems` (:136): the reference SOURCE is left alone, the parser's `_headers` OBJECT is
swapped for a dict subclass that has the Python-2 method.  The gzip branch (:125-128) feeds str to
zlib and cannot run here: gzip messages are left out (tests cover them by round trip).  Header
ORDER is Python-2 dict hash order in the reference and arrival order here (HttpParser.py:23):
consumers compare first line, body and the header lines as a multiset."""
import gzip
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import make_golden as G            # noqa: E402  (loads the reference through the cStringIO shim)


class Py2Dict(dict):
    def viewitems(self):
        return self.items()


def main():
    with gzip.open(os.path.join(HERE, 'parser_golden.json.gz')) as f:
        whole = json.loads(f.read().decode('ascii'))['whole']
    out = []
    for w in whole:
        obs = w['obs']
        if obs['exc'] or not obs['message_complete'] or 'gzip' in str(obs['headers'].get('content-encoding', '')):
            continue
        p = G.HttpParser()
        data = w['msg']
        while data and not p.message_complete():
            c = p.parse(data, len(data))
            data = data[c:]
        p._headers = Py2Dict(p._headers)
        out.append({'id': w['id'], 'str': str(p)})
    with gzip.GzipFile(os.path.join(HERE, 'serialize_golden.json.gz'), 'wb', mtime=0) as f:
        f.write(json.dumps(out, sort_keys=True).encode('ascii'))
    print('%d serialisations' % len(out))


if __name__ == '__main__':
    main()
