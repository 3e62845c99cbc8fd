import os
import sys
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def oracle():
    import oracle_py
    return oracle_py.load_oracle()


@pytest.fixture(scope="session")
def reference():
    import oracle_py
    if not oracle_py.reference_available():
        pytest.skip("oracle/_ref not built (needs /root/reference at build time)")
    return oracle_py.load_reference()


@pytest.fixture(scope="session")
def engine():
    import fgdp_b200
    if fgdp_b200.device_count() < 1:
        pytest.fail("no sm_100 GPU visible: -m gpu tests must run on the B200 box (there is no CPU fallback)")
    e = fgdp_b200.Engine()
    yield e
    e.close()
