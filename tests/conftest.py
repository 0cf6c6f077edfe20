"""pytest configuration: registers the `gpu` marker and shared scene fixtures.

Tests marked `gpu` run on a real B200 (`pytest -m gpu`); everything else runs on CPU.
/root/reference does not exist on the GPU box: scenes come from tests/golden/*.npz
(generated here by tests/golden/make_scene_fixtures.py) or from procedural meshes.
"""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

GOLDEN = os.path.join(ROOT, "tests", "golden")
REFERENCE = "/root/reference"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on the B200 box)")


def load_scene_fixture(name):
    """Six reference-layout blobs (uint8 arrays) stored by make_scene_fixtures.py."""
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    return {k: np.ascontiguousarray(z[k]) for k in ("spheres", "materials", "points", "triangles", "objects", "nodes")}


@pytest.fixture(scope="session")
def cornell_blobs():
    return load_scene_fixture("scene_cornell")


@pytest.fixture(scope="session")
def bunny_blobs():
    return load_scene_fixture("scene_bunny")


@pytest.fixture(scope="session")
def have_reference():
    return os.path.isdir(os.path.join(REFERENCE, "assets"))
