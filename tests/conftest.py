import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "opencl-raytracer_b200"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def oracle_mod():
    """The CPU oracles (test infrastructure).  Builds the C restatement if needed."""
    import subprocess

    import oracle

    if not os.path.exists(os.path.join(ROOT, "oracle", "liboracle_rt.so")):
        subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "port"])
    return oracle


@pytest.fixture(scope="session")
def rt():
    """The product library through its ctypes binding; fails loudly if it is not built."""
    from pyrt import lib

    lib.load()
    return lib


@pytest.fixture()
def ctx(rt):
    c = rt.Context(0)
    yield c
    c.close()
