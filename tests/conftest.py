import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def ctx_factory():
    """Creates mimi_b200 contexts; fails loudly (no skip, no fallback) when the CUDA path is unusable."""
    from mimi_b200 import capi

    made = []

    def make():
        c = capi.Context(0)
        made.append(c)
        return c

    yield make
    for c in made:
        c.close()
