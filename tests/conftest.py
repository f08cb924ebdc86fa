import importlib
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def _ensure_built():
    pkg = os.path.join(ROOT, "software-path-tracer_b200", "libcudapt.so")
    twin = os.path.join(ROOT, "oracle", "libtwin.so")
    ref = os.path.join(ROOT, "oracle", "_ref", "libref_render.so")
    need = not (os.path.exists(pkg) and os.path.exists(twin))
    if os.path.isdir("/root/reference/libs/render/src") and not os.path.exists(ref):
        need = True
    if need:
        subprocess.run(["make", "-C", ROOT, "all"], check=True, capture_output=True)


@pytest.fixture(scope="session")
def built():
    _ensure_built()
    return True


@pytest.fixture(scope="session")
def oracle(built):
    return importlib.import_module("oracle")


@pytest.fixture(scope="session")
def pt(built):
    """The product binding (ctypes over libcudapt.so)."""
    return importlib.import_module("software-path-tracer_b200")


@pytest.fixture(scope="session")
def scenes(built):
    return importlib.import_module("software-path-tracer_b200.scenes")


@pytest.fixture(scope="session")
def golden():
    import numpy as np
    return np.load(os.path.join(ROOT, "tests", "golden", "c1_spheres_96x54.npz"))


@pytest.fixture(scope="session")
def golden_mesh():
    """Twin-generated (tests/golden/make_golden_mesh.py): freezes the mesh specification; not pinned by the reference."""
    import numpy as np
    return np.load(os.path.join(ROOT, "tests", "golden", "c2_cornell_64x36.npz"))
