import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def ctx():
    from rna_clique_b200 import _native
    c = _native.Context(0)
    _native.debug_keep_hsps(c)   # tests compare the gapped-HSP set itself with the oracle's
    yield c
    c.close()
