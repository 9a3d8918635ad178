import os
import sys

import numpy as np
import pytest

# several contexts (5 streams each) share one GPU in the vr_group tests: with the default of 8 hardware queues their streams
# would alias and a spinning barrier kernel could block the peer it waits for; must be set before CUDA initialises
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _ensure_built():
    import __graft_entry__ as g
    lib = os.path.join(ROOT, "volumerenderer_b200", "libvr.so")
    orc = os.path.join(ROOT, "oracle", "libvr_oracle.so")
    if not os.path.exists(lib) or not os.path.exists(orc):
        g.build()


@pytest.fixture(scope="session")
def vr():
    _ensure_built()
    import volumerenderer_b200
    return volumerenderer_b200


@pytest.fixture(scope="session")
def pyoracle():
    _ensure_built()
    from oracle import pyoracle as m
    return m


@pytest.fixture(scope="session")
def oracle(pyoracle):
    return pyoracle.Oracle()


@pytest.fixture(scope="session")
def ref(pyoracle):
    """oracle/_ref: the reference's shader sources compiled as C++ (built where /root/reference exists)."""
    if not pyoracle.have_ref():
        pytest.skip("oracle/_ref/libvr_ref.so not built (needs the reference checkout)")
    return pyoracle.Ref()


_grids = {}


@pytest.fixture(scope="session")
def cloud(vr):
    """cloud(n): the SURVEY 8(d) procedural cloud (voxel 64/n, translation (-32,-12,-32)), cached per session."""
    def make(n, **kw):
        key = (n, tuple(sorted(kw.items())))
        if key not in _grids:
            _grids[key] = vr.build_cloud_grid(n, **kw)
        return _grids[key]
    return make


@pytest.fixture(scope="session")
def has_gpu():
    import ctypes
    try:
        cuda = ctypes.CDLL("libcudart.so")
    except OSError:
        try:
            import torch
            return torch.cuda.is_available()
        except Exception:
            return False
    n = ctypes.c_int(0)
    return cuda.cudaGetDeviceCount(ctypes.byref(n)) == 0 and n.value > 0


def rel_err(a, b):
    """max relative error with the |ref| < 1e-6 -> absolute rule of BASELINE.json / SURVEY 8(d), and RMSE."""
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    denom = np.where(np.abs(b) < 1e-6, 1.0, np.abs(b))
    return float(np.max(np.abs(a - b) / denom)), float(np.sqrt(np.mean((a - b) ** 2)))
