import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def ftf():
    """The product package (loads libftf_b200.so; building it first if this checkout has not)."""
    import __graft_entry__
    if not os.path.exists(os.path.join(ROOT, "flink-tensorflow_b200", "libftf_b200.so")):
        __graft_entry__.build()
    import ftf_loader
    return ftf_loader.load()


@pytest.fixture(scope="session")
def oracle():
    import oracle as o
    o.lib()
    return o


@pytest.fixture(scope="session")
def golden():
    import json
    out = {}
    gdir = os.path.join(ROOT, "tests", "golden")
    for name in os.listdir(gdir):
        if name.endswith(".json"):
            with open(os.path.join(gdir, name)) as f:
                out[name[:-5]] = json.load(f)
    return out


@pytest.fixture(scope="session")
def model_dir():
    return os.path.join(ROOT, "models", "half_plus_two")
