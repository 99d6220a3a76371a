"""pytest configuration: `gpu` marker (tests that need a B200) and shared fixtures."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "direct-visual-odometry_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def orc():
    from oracle.oracle import Oracle
    return Oracle()


@pytest.fixture(scope="session")
def dvolib():
    import dvo_b200
    return dvo_b200.load_library()


@pytest.fixture(scope="session")
def seq160():
    """6 frames of a synthetic 640x480 sequence + GT depth (main.cpp geometry: levels 3, culls 2)."""
    from dvo_b200 import synth
    return synth.sequence(6, 640, 480, seed=1234, trans_step=0.006, rot_step_deg=0.25)


def make_dvo(**kw):
    import dvo_b200
    return dvo_b200.Dvo(**kw)
