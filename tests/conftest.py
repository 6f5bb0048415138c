import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "gaussian-transformer_b200")
for p in (PKG, os.path.dirname(os.path.abspath(__file__)), ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def oracle():
    from checkers import Oracle
    return Oracle()


@pytest.fixture(scope="session")
def reference():
    from checkers import Reference, REF_SO
    if not os.path.exists(REF_SO):
        pytest.skip("oracle/_ref/libref_rasterizer.so not built (needs /root/reference at build time)")
    return Reference()
