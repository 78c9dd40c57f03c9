import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "path-tracing_b200"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, os.path.join(ROOT, "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    # the oracle's C restatement is test infrastructure: build it if missing (gcc only, a second or two)
    if not os.path.exists(os.path.join(ROOT, "oracle", "libptoracle.so")):
        subprocess.run(["make", "-C", os.path.join(ROOT, "oracle")], check=False, capture_output=True)


@pytest.fixture(scope="session")
def oracle():
    from oracle_py import Oracle
    return Oracle()


@pytest.fixture(scope="session")
def reference():
    """The unmodified reference compiled with the shims (oracle/_ref); absent sources -> skip."""
    from oracle_py import Reference
    if not Reference.available():
        pytest.skip("oracle/_ref/libptref.so not built (needs /root/reference)")
    return Reference()


@pytest.fixture(scope="session")
def lib():
    import ptb200
    return ptb200.Library()


@pytest.fixture(scope="session")
def loader():
    import ptb200
    return ptb200.HostLoader()


@pytest.fixture(scope="session")
def tor(loader):
    """models/Tor.obj through the product's C++ loader."""
    return loader.load_obj(os.path.join(ROOT, "assets", "models"), "Tor.obj")
