import json
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def I(x):
    """decimal-string fixtures -> ints (recursively; lists become tuples)."""
    if isinstance(x, (list, tuple)):
        return tuple(I(y) for y in x)
    return int(x)


_cache = {}


def load_golden(name):
    if name not in _cache:
        with open(os.path.join(GOLDEN, name)) as fh:
            _cache[name] = json.load(fh)
    return _cache[name]


@pytest.fixture(params=["tiny", "t0", "node", "t1"])
def tier(request):
    return load_golden("ref_%s.json" % request.param)
