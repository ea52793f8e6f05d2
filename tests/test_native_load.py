"""CPU-side checks of the C-ABI boundary: the library builds/loads, exports every symbol that
include/shellac_b200.h declares, and refuses to work (loudly) without a CUDA device."""
import ctypes as C
import os

import pytest

from shellac_b200 import _native as N
from shellac_b200 import build as B


@pytest.fixture(scope='module')
def lib():
    B.build()
    return N.lib()


def test_library_exports_every_declared_symbol(lib):
    declared = N.exported_symbols()
    assert len(declared) >= 35
    raw = C.CDLL(N.LIB_PATH)
    missing = [s for s in declared if not hasattr(raw, s)]
    assert not missing, missing


def test_struct_layouts_match_header():
    assert N.HDR_DTYPE.itemsize == 8 and N.CHUNK_DTYPE.itemsize == 16
    assert C.sizeof(N.ParseOutStruct) == 18 * 8
    assert C.sizeof(N.FrameOutStruct) == 9 * 8
    assert C.sizeof(N.PlaceOutStruct) == 4 * 8


def test_no_cpu_fallback_without_device(lib):
    if lib.shl_device_count() > 0:
        pytest.skip('a CUDA device is present')
    from shellac_b200.engine import Engine
    with pytest.raises(N.ShellacError):
        Engine()
    h = C.c_void_p()
    assert lib.shl_ctx_create(0, 16, 1024, 0, C.byref(h)) == -4      # SHL_ERR_NO_DEVICE
    assert b'no CPU fallback' in lib.shl_last_error(None)


def test_host_corpus_is_deterministic_and_well_formed(lib):
    """The counter-based generator (configs 2/4/5) on the host: same bytes for the same (seed, i),
    lengths agree with the fill, and the oracle parses every message as a complete GET."""
    import numpy as np
    from shellac_b200.batch import Arena
    from tests import oracle_lib as O
    n = 2000
    ln = np.zeros(n, np.uint32)
    assert lib.shl_corpus_lengths_host(20130725, 0, n, 0, ln.ctypes.data) == 0
    off = np.zeros(n, np.uint64)
    np.cumsum(ln[:-1], dtype=np.uint64, out=off[1:])
    data = np.zeros(int(off[-1] + ln[-1]) + 64, np.uint8)
    assert lib.shl_corpus_fill_host(20130725, 0, n, 0, off.ctypes.data, data.ctypes.data) == 0
    d2 = np.zeros_like(data)
    lib.shl_corpus_fill_host(20130725, 0, n, 0, off.ctypes.data, d2.ctypes.data)
    assert (data == d2).all()
    assert 256 <= ln.min() and ln.max() <= 1100 and 480 < ln.mean() < 545
    ar = Arena(data, off, ln)
    po = O.orc_parse(ar)
    assert (po.code() == N.S_OK).all()
    assert ((po.status[:n] & N.F_MESSAGE_COMPLETE) != 0).all()
    assert 8 <= po.n_headers[:n].min() and po.n_headers[:n].max() <= 20
    assert (po.first_line_len[:n] + po.headers_len[:n] == ln).all()
    # sliding the window reproduces the same messages (message i is a function of (seed, i) only)
    ln2 = np.zeros(10, np.uint32)
    lib.shl_corpus_lengths_host(20130725, 500, 10, 0, ln2.ctypes.data)
    assert (ln2 == ln[500:510]).all()
