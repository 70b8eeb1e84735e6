import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def port():
    from oracle import oracle_lib
    return oracle_lib.load("port")


@pytest.fixture(scope="session")
def ref():
    """The compiled reference (oracle/_ref/libpv3_ref.so); skipped where it has not been built."""
    from oracle import oracle_lib
    if not oracle_lib.have("reference"):
        pytest.skip("oracle/_ref/libpv3_ref.so not built (needs /root/reference: make -C oracle ref)")
    return oracle_lib.load("reference")


@pytest.fixture(scope="session")
def fixtures_json():
    import json
    with open(os.path.join(GOLDEN, "reference_fixtures.json")) as fh:
        return json.load(fh)


@pytest.fixture(scope="session")
def workflow_msa():
    from panvc3_b200.workload import load_msa_npz
    return load_msa_npz(os.path.join(GOLDEN, "test_workflow_msa.npz"))
