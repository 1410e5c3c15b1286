import os
import sys

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
for p in (ROOT, HERE):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def oracle():
    import oracle_lib
    return oracle_lib.oracle()


@pytest.fixture(scope="session")
def ref():
    """The reference's own code compiled into oracle/_ref (None -> tests that need it skip)."""
    import oracle_lib
    lib = oracle_lib.ref()
    if lib is None:
        pytest.skip("oracle/_ref not built (needs /root/reference at build time)")
    return lib
