import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")
FIX = os.path.join(GOLDEN, "fixtures")
MATERIAL = os.path.join(FIX, "materials", "B9Creator.material")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    config.addinivalue_line("markers", "slow: long-running pin (minutes)")


def bc_path(name):
    return os.path.join(FIX, "bcs", name)


@pytest.fixture(scope="session")
def oracle_mod():
    from oracle import oracle as O
    O.build()
    return O


def pytest_collection_modifyitems(config, items):
    """The multi-rank-on-one-GPU slab tests run last: they are the longest and the only ones that drive several host
    threads, so `-x` still gets through the single-rank parity tests first.
    tests/test_zz_no_alloc.py (an entry point outside the reference's surface) stays the very last."""
    items.sort(key=lambda it: 2 if "test_zz_no_alloc" in it.nodeid else 1 if "test_gpu_slab" in it.nodeid else 0)
