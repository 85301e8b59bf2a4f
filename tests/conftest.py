import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def port():
    import oracle
    oracle.build(("port",))
    return oracle.PortOracle()


@pytest.fixture(scope="session")
def ref():
    """The compiled reference (oracle/_ref).  Present wherever /root/reference was available at build time;
    the prebuilt .so travels to the GPU box."""
    import oracle
    if not oracle.ref_available():
        if os.path.isdir("/root/reference"):
            oracle.build(("ref",))
        else:
            pytest.skip("oracle/_ref not built and /root/reference absent")
    return oracle.RefOracle()


@pytest.fixture(scope="session")
def checker(port):
    """Best available checker: the compiled reference, else the port."""
    import oracle
    return oracle.best_oracle()


@pytest.fixture(scope="session")
def lib():
    import cilqr_b200
    from cilqr_b200 import solver
    return solver.load_library()
