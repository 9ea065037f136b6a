import os
import sys

import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "madrona-mp-env_b200"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def built():
    """Native artefacts the CPU suite needs: oracle (prebuilt on the GPU box) and hostsim."""
    import simlib
    simlib.build_oracle()
    simlib.build_hostsim()
    return simlib
