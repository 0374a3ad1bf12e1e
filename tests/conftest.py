import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real CUDA device (run with -m gpu on a B200)")


@pytest.fixture(scope="session", autouse=True)
def _native_built():
    """Build the native libraries once if a fresh checkout has none."""
    lib = os.path.join(ROOT, "b200quad", "lib", "libb200quad.so")
    cpu = os.path.join(ROOT, "b200quad", "lib", "libb200quad_cpu.so")
    if not (os.path.exists(lib) and os.path.exists(cpu)):
        import build
        build.build_core()
        build.build_torch()
        build.build_cli()
    yield


def gpu_count():
    from b200quad import _native
    return _native.cuda_device_count()
