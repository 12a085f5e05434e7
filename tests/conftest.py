import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def oracle():
    from oracle_lib import Oracle
    return Oracle()


@pytest.fixture(scope="session")
def ref():
    from oracle_lib import Ref, have_ref
    if not have_ref():
        pytest.skip("oracle/_ref not built (needs /root/reference at build time)")
    r = Ref()
    r.set_threads(min(8, os.cpu_count() or 1))
    return r


@pytest.fixture(scope="session")
def gm():
    """libgm_cuda on cuda:0 -- fails loudly (no skip, no fallback) when the library or the GPU is missing."""
    from gmb200_loader import load_package
    pkg = load_package()
    lib = pkg.GmCuda()
    lib.init(0)
    return lib
