import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "tools"), os.path.join(ROOT, "rgbd-slam_b200", "py")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def oracle():
    import orclib

    orclib.lib()
    return orclib


@pytest.fixture(scope="session")
def seq_vga():
    """Four frames of the TUM-fr1-shaped synthetic sequence (config 1 shape)."""
    import synth

    return synth.sequence("tum_fr1", n=4, seed=1234)
