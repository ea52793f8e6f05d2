"""N2, message re-serialisation (HttpParser.__str__, HttpParser.py:111-138; callers Server.py:361,
418): the drop-in class's serialize() / str() and the CPython port against str(parser) of the
UNMODIFIED reference over every complete, exception-free message of the golden corpus
(tests/golden/serialize_golden.json.gz, made by tests/golden/make_serialize_golden.py).  Header
order is Python-2 dict order in the reference: first line, body and the header lines as a multiset."""
import gzip
import json
import os
import sys
import zlib

import pytest

from shellac_b200.server import HttpParser
from tests import oracle_lib as O
from tests.oracle_backend import OracleBackend

sys.path.insert(0, os.path.join(O.ROOT, 'oracle'))


def _golden_str():
    with gzip.open(os.path.join(O.ROOT, 'tests', 'golden', 'serialize_golden.json.gz')) as f:
        return json.loads(f.read().decode('ascii'))


def _split(s):
    head, _, body = s.partition('\r\n\r\n')
    lines = head.split('\r\n')
    return lines[0], sorted(lines[1:]), body


def _check_backend(be, golden):
    by_id = {w['id']: w['msg'] for w in golden['whole']}
    want = _golden_str()
    ps = HttpParser.parse_many([by_id[g['id']] for g in want], backend=be)
    n_hdr = 0
    for g, p in zip(want, ps):
        assert p.error is None and p.message_complete(), g['id']
        assert _split(str(p)) == _split(g['str']), g['id']
        n_hdr += len(_split(g['str'])[1])
    assert len(want) > 2000 and n_hdr > 5000


def test_facade_serialize_matches_reference(golden):
    _check_backend(OracleBackend(), golden)


@pytest.mark.gpu
def test_facade_serialize_matches_reference_on_device(golden):
    from shellac_b200.server.HttpParser import DeviceBackend
    _check_backend(DeviceBackend(), golden)


def test_port_serialize_matches_reference(golden):
    import py_port
    by_id = {w['id']: w['msg'] for w in golden['whole']}
    for g in _golden_str():
        p, calls, exc = py_port.drive(by_id[g['id']].encode('latin-1'))
        assert exc is None
        assert _split(py_port.serialize(p).decode('latin-1')) == _split(g['str']), g['id']


def test_serialize_gzip_round_trip(golden):
    """The gzip branch (:125-128) cannot run in the Python-3-hosted reference: the re-serialised body
    must inflate back to the plaintext the reference's own test asserts (HttpParserTests.py:247-252)
    and content-length must be recomputed from the compressed bytes."""
    for name, z in golden['gzip'].items():
        p = HttpParser.parse_many([z['msg']], backend=OracleBackend())[0]
        first, hdrs, _ = _split(str(p))
        raw = p.serialize()
        body = raw[raw.index(b'\r\n\r\n') + 4:]
        assert zlib.decompress(body, 31).decode('latin-1') == z['plain'], name
        assert 'Content-Length: %d' % len(body) in hdrs and not any(h.lower().startswith('transfer-encoding') for h in hdrs)
