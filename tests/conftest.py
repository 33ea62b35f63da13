import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def oracle_built():
    """The oracle port is compiled on demand (gcc, a second or two)."""
    from oracle import binding
    if not os.path.exists(binding.PORT_SO):
        binding.build_port()
    return binding


@pytest.fixture(scope="session")
def ref_built(oracle_built):
    """The unmodified reference, compiled from /root/reference when that tree is mounted;
    on the GPU box the prebuilt oracle/_ref travels with the snapshot."""
    b = oracle_built
    if not b.have_ref():
        if not b.build_ref():
            pytest.skip("reference tree not mounted and oracle/_ref not prebuilt")
    return b
