"""CPU tests: the C-ABI library loads, exports every symbol include/b200reg.h declares, and fails loudly without a GPU."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    with open(os.path.join(ROOT, "include", "b200reg.h")) as f:
        text = f.read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b200reg_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_are_exported_and_bound():
    from open3d_slam_advanced_rss_2024_public_b200 import capi
    syms = declared_symbols()
    assert len(syms) >= 35
    L = ctypes.CDLL(capi.LIB_PATH)
    for s in syms:
        assert hasattr(L, s), f"libb200reg.so does not export {s}"
    assert sorted(capi.SYMBOLS) == syms, set(capi.SYMBOLS) ^ set(syms)
    assert capi.lib().b200reg_abi_version() == 2


def test_no_cpu_fallback_without_gpu():
    from open3d_slam_advanced_rss_2024_public_b200 import capi
    try:
        n = ctypes.c_int(0)
        has_gpu = ctypes.CDLL("libcudart.so").cudaGetDeviceCount(ctypes.byref(n)) == 0 and n.value > 0
    except OSError:
        has_gpu = False
    if has_gpu:
        pytest.skip("a GPU is present")
    with pytest.raises(capi.CudaError):
        capi.Context(0)


def test_product_does_not_import_the_oracle():
    pkg = os.path.join(ROOT, "open3d_slam_advanced_rss_2024_public_b200")
    for dirpath, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".cpp", ".h", ".hpp")):
                with open(os.path.join(dirpath, fn)) as f:
                    src = f.read()
                assert "oracle_py" not in src and "liboracle" not in src and "oracle.h" not in src, fn


def test_solve_point_to_plane_host_side(oracle):
    """b200reg_solve_point_to_plane is pure host arithmetic: checkable without a GPU against the oracle's solver."""
    import numpy as np
    from open3d_slam_advanced_rss_2024_public_b200 import capi
    rng = np.random.default_rng(0)
    n = 2000
    ref = rng.normal(size=(n, 3)).astype(np.float32) * 2
    nrm = rng.normal(size=(n, 3)); nrm = (nrm / np.linalg.norm(nrm, axis=1, keepdims=True)).astype(np.float32)
    reading = (ref + rng.normal(size=(n, 3)) * 0.01 + np.array([0.02, -0.01, 0.03])).astype(np.float32)
    ids = np.arange(n, dtype=np.int32).reshape(n, 1)
    d2 = np.zeros((n, 1), dtype=np.float32); w = np.ones((n, 1), dtype=np.float32)
    for f4 in (False, True):
        o = oracle.point_to_plane(oracle.feat4(reading), oracle.feat4(ref), nrm, ids, d2, w, force4DOF=f4, precision=1)
        o6 = oracle.point_to_plane(oracle.feat4(reading), oracle.feat4(ref), nrm, ids, d2, w, precision=1)
        T, x = capi.solve_point_to_plane(o6["A"], o6["b"], reading.astype(np.float64).mean(0), ref.astype(np.float64).mean(0), force4DOF=f4)
        assert np.abs(T - o["T"]).max() < 1e-6
    # rank-deficient system (planar grid, icpSingular): minimum-norm solution = pure z translation
    xs = np.array([[0.1 * x - 0.5, 0.1 * y - 0.5, 0.0] for x in range(10) for y in range(10)], dtype=np.float32)
    rf = xs.copy(); rf[:, 2] = 1.0
    nz = np.tile(np.array([[0, 0, 1]], dtype=np.float32), (100, 1))
    ids = np.arange(100, dtype=np.int32).reshape(100, 1)
    o = oracle.point_to_plane(oracle.feat4(xs), oracle.feat4(rf), nz, ids, np.ones((100, 1), np.float32), np.ones((100, 1), np.float32), precision=1)
    T, x = capi.solve_point_to_plane(o["A"], o["b"], xs.astype(np.float64).mean(0), rf.astype(np.float64).mean(0))
    E = np.eye(4); E[2, 3] = 1
    assert np.allclose(T, E, atol=1e-6) and np.allclose(o["T"], E, atol=1e-6)


def test_mapper_driver_symbols_are_exported():
    """libb200pm.so carries the open3d_slam Mapper driver (host/mapper.h) behind b200map_* (no compute without a GPU)."""
    from open3d_slam_advanced_rss_2024_public_b200 import mapper
    L = mapper._lib()
    for name in ("b200map_create", "b200map_destroy", "b200map_set_pose", "b200map_add_range_measurement", "b200map_add_pointcloud2", "b200map_prefetch", "b200map_map"):
        assert hasattr(L, name)
