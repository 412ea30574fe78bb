import os
import sys

import pytest

# The host-buffer entry points cut a batch into pipeline chunks; the smallest chunk size puts the most chunk
# boundaries (and both staging slots, several times over) inside the test batches.  The library reads this once.
os.environ.setdefault("NBG_CHUNK_SCALE", "1")

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def ctx():
    """One nbg context on cuda:0 for the whole GPU session (fails loudly without a GPU)."""
    import importlib
    pkg = importlib.import_module("odh-kubeflow_b200")
    c = pkg.Context([0])
    yield c
    c.close()
