import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def _has_gpu():
    try:
        import ctypes
        cudart = ctypes.CDLL("libcudart.so")
        n = ctypes.c_int(0)
        return cudart.cudaGetDeviceCount(ctypes.byref(n)) == 0 and n.value > 0
    except OSError:
        try:
            import torch
            return torch.cuda.is_available()
        except Exception:
            return False


def pytest_collection_modifyitems(config, items):
    # GPU tests are selected with -m gpu on the GPU box; without a device they are skipped, never silently passed.
    if _has_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container (the product has no CPU fallback)")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def golden():
    g = os.path.join(ROOT, "tests", "golden")
    car = np.load(os.path.join(g, "car_clouds.npz"))
    vtk = np.load(os.path.join(g, "vtk_clouds.npz"))
    with open(os.path.join(g, "golden_transforms.json")) as f:
        T = {k: np.array(v) for k, v in json.load(f).items()}
    return dict(car=car, vtk=vtk, T=T)


@pytest.fixture(scope="session")
def oracle():
    from oracle import oracle_py
    oracle_py.lib()
    return oracle_py


@pytest.fixture(scope="session")
def ctx():
    from open3d_slam_advanced_rss_2024_public_b200 import capi
    c = capi.Context(0)
    yield c
    c.close()
