import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def oracle():
    """The compiled reference (oracle/_ref).  Built here when /root/reference is mounted;
    on the GPU box the prebuilt libraries travel with the snapshot."""
    from oracle import refharness
    if not refharness.available("parity"):
        if os.path.isdir("/root/reference/deps/opende"):
            refharness.build("parity")
        else:
            pytest.skip("oracle/_ref not built and /root/reference not mounted")
    return refharness


@pytest.fixture(scope="session")
def gzlib():
    from gazebo_b200 import build, batch
    if not os.path.exists(batch.LIB_PATH):
        build.build_lib()
    return batch.lib()
