import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def orc():
    """The CPU oracle (test infrastructure, oracle/cpu_odom.cpp)."""
    from oracle import oracle
    oracle.build()
    return oracle


@pytest.fixture(scope="session")
def synth():
    from densemonoslam_b200 import synth as s
    return s


def small_cam(synth, div=4):
    return synth.scaled_camera(synth.CAM_VGA, div)


@pytest.fixture(scope="session")
def pair_small(synth):
    cam = small_cam(synth, 4)  # 160x120
    model, live, T = synth.frame_pair(cam, seed=7)
    return cam, model, live, T


@pytest.fixture(scope="session")
def pair_vga(synth):
    cam = synth.CAM_VGA
    model, live, T = synth.frame_pair(cam, seed=1234)
    return cam, model, live, T


def assert_maps_equal(a, b, what=""):
    """Planar maps [3,H,W]: x planes identical (incl. NaN pattern); y/z compared where x is valid."""
    ax, bx = a[0], b[0]
    assert np.array_equal(np.isnan(ax), np.isnan(bx)), f"{what}: NaN pattern differs"
    ok = ~np.isnan(ax)
    for p in range(3):
        assert np.array_equal(a[p][ok].view(np.uint32), b[p][ok].view(np.uint32)), f"{what}: plane {p} differs"


def assert_f32_equal(a, b, what=""):
    a = np.asarray(a, np.float32); b = np.asarray(b, np.float32)
    assert a.shape == b.shape, what
    same = (a.view(np.uint32) == b.view(np.uint32)) | (np.isnan(a) & np.isnan(b))
    assert same.all(), f"{what}: {np.count_nonzero(~same)} of {a.size} values differ, max abs {np.nanmax(np.abs(a - b))}"
