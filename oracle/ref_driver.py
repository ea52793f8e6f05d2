"""oracle/ref_driver.py — TEST / BENCH INFRASTRUCTURE: the reference's CPU path, as bench.py times it.

`load_reference()` imports the UNMODIFIED reference parser from the bytecode oracle/build_ref.py
made (oracle/_ref/HttpParser.bc), behind the same three-line `cStringIO` shim SURVEY.md section 8c
describes (the reference is Python 2; under CPython 3 it runs on `str`, one char per byte).  When
oracle/_ref is absent the CPython port (oracle/py_port.py, pinned to the same golden vectors) stands in.

`route_reference(msgs, ring)` is the path of Server._read_requests for a batch of whole requests
(/root/reference/src/python/shellac/server/Server.py:320-335): drive HttpParser.parse over each
message like test/HttpParserTests.py:36-39, key = request.url(), and place the key.  Placement in
the reference is pylibmc -> libmemcached (third-party C, absent here): the hashlib + bisect
restatement of py_port stands in for it (PARITY UNPINNED).
"""
import importlib.machinery
import importlib.util
import io
import os
import sys
import types

HERE = os.path.dirname(os.path.abspath(__file__))
if HERE not in sys.path:
    sys.path.insert(0, HERE)
import py_port  # noqa: E402

REF = os.path.join(HERE, '_ref')


def load_reference():
    """(HttpParser class, 'reference') from oracle/_ref, or (None, 'port')."""
    pyc = os.path.join(REF, 'HttpParser.bc')
    if not os.path.exists(pyc):
        print('oracle/_ref/HttpParser.bc absent: the CPython port stands in', file=sys.stderr)
        return None, 'port'
    if 'cStringIO' not in sys.modules:
        m = types.ModuleType('cStringIO')
        m.StringIO = io.StringIO
        sys.modules['cStringIO'] = m
    try:
        loader = importlib.machinery.SourcelessFileLoader('HttpParser', pyc)
        spec = importlib.util.spec_from_loader('HttpParser', loader)
        mod = importlib.util.module_from_spec(spec)
        loader.exec_module(mod)
        return mod.HttpParser, 'reference'
    except Exception as e:  # noqa: BLE001 - bytecode of another interpreter version
        print('oracle/_ref present but not loadable (%r): the CPython port stands in' % (e,), file=sys.stderr)
        return None, 'port'


def route_reference(msgs, values, nodes, n_hosts, parser_cls):
    """msgs: list of latin-1 `str`.  Returns (complete requests, per-node histogram)."""
    place = py_port.place
    hist = [0] * n_hosts
    done = 0
    for data in msgs:
        p = parser_cls()
        while data and not p.message_complete():
            c = p.parse(data, len(data))
            data = data[c:]
        if p.message_complete():
            key = p.url()
            hist[place(values, nodes, n_hosts, key.encode('latin-1'))] += 1
            done += 1
    return done, hist


def place_keys(keys, values, nodes, n_hosts):
    """Config 5's CPU side: the placement restatement alone over `bytes` keys."""
    place = py_port.place
    return [place(values, nodes, n_hosts, k) for k in keys]
