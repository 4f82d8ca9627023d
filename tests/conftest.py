import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")
    # Build what is missing (the GPU box receives prebuilt .so files; this is a no-op there).
    need = [os.path.join(ROOT, "n-body-simulation_b200", "libnbody_b200.so"),
            os.path.join(ROOT, "oracle", "libnboracle.so")]
    if not all(os.path.exists(p) for p in need):
        subprocess.run(["make", "-C", ROOT, "lib", "oracle", "ref"], check=False,
                       stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)


@pytest.fixture(scope="session")
def oracle():
    import nbtest
    return nbtest.Oracle()


@pytest.fixture(scope="session")
def ref():
    import nbtest
    r = nbtest.Ref.try_load()
    if r is None:
        pytest.skip("oracle/_ref/libnbref.so not built (reference tree absent)")
    return r


@pytest.fixture(scope="session")
def lib():
    import nbtest
    return nbtest.Lib()
