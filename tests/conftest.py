import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line('markers', 'gpu: needs a CUDA device (run with -m gpu on a B200)')


@pytest.fixture(scope='session')
def golden():
    from tests.oracle_lib import load_golden
    return load_golden()


@pytest.fixture(scope='session')
def placement_golden():
    import json
    with open(os.path.join(ROOT, 'tests', 'golden', 'placement_golden.json')) as f:
        return json.load(f)
