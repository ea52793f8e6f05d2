#!/usr/bin/env python
"""Run GPU parity tests against the -DSHL_BOUNDS_CHECK build (tools/build_variants.py chk:-DSHL_BOUNDS_CHECK)
and report how many out-of-bounds indices the kernels counted.  compute-sanitizer is closed on this
pool, so these checks of our own stand in for memcheck on the indices that matter (shared-memory
bitmaps, record slots).   SHL_LIB=build/variants/lib_chk.so python tools/bounds_check_run.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pytest                                    # noqa: E402
from shellac_b200 import _native as N            # noqa: E402

rc = pytest.main(['tests/test_gpu_parity.py', '-m', 'gpu', '-q', '-x', '-k',
                  'golden or structural or token_soup or unordered or c2_corpus or route_fused or lowercased or key_too_long or empty_and_tiny'])
print('pytest rc=%d  shl_debug_oob_count=%d  (library %s)' % (rc, N.lib().shl_debug_oob_count(), N.LIB_PATH))
sys.exit(rc)
