# SPDX-License-Identifier: WTFPL
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def oracle():
    import oracle_harness
    return oracle_harness.Oracle()


@pytest.fixture(scope="session")
def reference():
    """The reference's own shader code compiled as C++ (oracle/_ref); absent where /root/reference never was."""
    import oracle_harness
    ref = oracle_harness.Reference.open()
    if ref is None:
        pytest.skip("oracle/_ref is not built (no /root/reference on this machine)")
    return ref


@pytest.fixture(scope="session")
def vcrt():
    """The product: ctypes view of lib/libvcrt.so. Fails (does not skip) when the library is missing."""
    from vulkancomputeraytracing_b200 import cabi
    return cabi.load()


@pytest.fixture()
def gpu_scene(vcrt):
    """Reference scene uploaded to cuda:0; tears the renderer down afterwards."""
    from vulkancomputeraytracing_b200 import host
    scene = host.generate_scene()
    host.upload_scene(scene)
    yield scene
    host.destroy()
