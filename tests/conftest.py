import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def sdr():
    import sdrb200_pkg
    import testkit
    return testkit.attach(sdrb200_pkg.load())


@pytest.fixture(scope="session")
def oracle():
    """the reference build when present, else the CPU port (oracle/)."""
    import oracle_bind as ob
    if not (ob.have_ref() or ob.have_port()):
        pytest.skip("no oracle library built (run __graft_entry__.build())")
    return ob.best()


@pytest.fixture(scope="session")
def oracles():
    import oracle_bind as ob
    out = []
    if ob.have_ref():
        out.append(ob.ref())
    if ob.have_port():
        out.append(ob.port())
    if not out:
        pytest.skip("no oracle library built")
    return out
