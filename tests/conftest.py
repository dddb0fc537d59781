import json
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def load_golden():
    with open(os.path.join(ROOT, "tests", "golden", "xet_golden.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def golden():
    return load_golden()


@pytest.fixture(scope="session")
def hasher():
    """One GPU context for the whole session (gpu-marked tests only)."""
    import ome_b200
    h = ome_b200.XetHasher(stage_timing=True)
    yield h
    h.close()
