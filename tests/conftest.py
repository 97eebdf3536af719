import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


def pytest_collection_modifyitems(config, items):
    import torch
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session")
def o3r():
    from oracle import o3r as mod
    mod.lib()
    return mod


@pytest.fixture(scope="session")
def small_kinect():
    """A reduced-resolution Kinect-shaped source (fast for the CPU oracle): 160x144 depth, 320x180 colour."""
    from azurekinect3dreconstruction_b200 import synthetic as syn
    d = dict(width=160, height=144, fx=504.2 / 4, fy=504.1 / 4, cx=319.8 / 4, cy=287.9 / 4)
    c = dict(width=320, height=180, fx=605.6 / 4, fy=605.9 / 4, cx=637.7 / 4, cy=364.3 / 4)
    return syn.SyntheticKinect(depth_intrinsics=d, color_intrinsics=c, seed=0, dropout=0.02)


def K_of(intr):
    return (intr["fx"], intr["fy"], intr["cx"], intr["cy"])
