import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def oracle():
    from oracle_lib import Oracle
    return Oracle()


@pytest.fixture(scope="session")
def reference():
    from oracle_lib import Reference, have_reference
    if not have_reference():
        pytest.skip("oracle/_ref/libmbgref.so not built (needs /root/reference; `make -C oracle ref`)")
    return Reference()


@pytest.fixture(scope="session")
def emu_library():
    """GPU-less emulator build of the kernels (tests/emu); debugging aid, not the product."""
    import subprocess
    d = os.path.join(ROOT, "tests", "emu")
    subprocess.check_call(["make", "-C", d], stdout=subprocess.DEVNULL)
    return os.path.join(d, "libmbgfe_emu.so")
