import json
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on the B200 box)")


@pytest.fixture(scope="session")
def scan_golden():
    with open(os.path.join(GOLDEN, "scan_golden.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def join_golden():
    with open(os.path.join(GOLDEN, "join_golden.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def minus_golden():
    with open(os.path.join(GOLDEN, "minus_golden.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def agg_golden():
    with open(os.path.join(GOLDEN, "agg_golden.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def pattern_golden():
    with open(os.path.join(GOLDEN, "pattern_golden.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def values_golden():
    with open(os.path.join(GOLDEN, "values_golden.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def bgp_golden():
    with open(os.path.join(GOLDEN, "bgp_golden.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def oracle():
    from oracle import oracle_py
    oracle_py.build()
    return oracle_py


@pytest.fixture(scope="session")
def ctx():
    """The product context.  Fails (does not skip) when the CUDA library or a GPU is missing: -m gpu tests must
    never pass on a silent fallback."""
    import rdf_fusion_b200 as rf
    c = rf.Context(0)
    yield c
    c.close()
