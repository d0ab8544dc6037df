"""The drop-in boundary: libpointmatcher's registrable-module surface (Registrar + Parametrizable + YAML chain loading)
mirrored by host/pm_compat.h, with the Cuda* modules registered by name.  CPU tests cover loading / error behaviour (no GPU
call is made before compute); GPU tests run the reference's own chain files with the module names swapped to their CUDA twins
and compare (a) module-by-module through the Eigen-shaped interfaces, (b) the fused fast lane, (c) the CPU oracle."""
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
YDIR = os.path.join(ROOT, "tests", "golden", "yaml")

CUDA_NAMES = {"KDTreeMatcher": "CudaKDTreeMatcher", "TrimmedDistOutlierFilter": "CudaTrimmedDistOutlierFilter",
              "SurfaceNormalOutlierFilter": "CudaSurfaceNormalOutlierFilter", "PointToPlaneErrorMinimizer": "CudaPointToPlaneErrorMinimizer",
              "VoxelGridDataPointsFilter": "CudaVoxelGridDataPointsFilter", "SurfaceNormalDataPointsFilter": "CudaSurfaceNormalDataPointsFilter"}


def yaml_text(name):
    with open(os.path.join(YDIR, name)) as f:
        return f.read()


def to_cuda(text):
    """What a user does to switch a chain file to the CUDA modules: prefix the module names."""
    for cpu, gpu in CUDA_NAMES.items():
        text = re.sub(r"(?<![A-Za-z])" + cpu + r"(?![A-Za-z])", gpu, text)
    return text


@pytest.fixture(scope="module")
def shim():
    from open3d_slam_advanced_rss_2024_public_b200 import pm_shim
    pm_shim.lib()
    return pm_shim


# ---------------------------------------------------------------------------------------------------------------
# CPU: registrar, parameter tables, YAML loading, error behaviour
# ---------------------------------------------------------------------------------------------------------------
def test_cuda_modules_are_registered_by_name(shim):
    assert {"CudaVoxelGridDataPointsFilter", "CudaSurfaceNormalDataPointsFilter"} <= set(shim.list_modules("DataPointsFilter"))
    assert "CudaKDTreeMatcher" in shim.list_modules("Matcher")
    assert {"CudaTrimmedDistOutlierFilter", "CudaSurfaceNormalOutlierFilter", "CudaMaxDistOutlierFilter", "CudaMinDistOutlierFilter",
            "CudaMedianDistOutlierFilter"} <= set(shim.list_modules("OutlierFilter"))
    assert {"CudaPointToPlaneErrorMinimizer", "CudaPointToPlaneWithCovErrorMinimizer", "CudaPointToPointErrorMinimizer"} <= set(shim.list_modules("ErrorMinimizer"))
    assert "CudaVarTrimmedDistOutlierFilter" in shim.list_modules("OutlierFilter")
    assert {"CudaMaxDistDataPointsFilter", "CudaMinDistDataPointsFilter", "CudaBoundingBoxDataPointsFilter",
            "CudaRandomSamplingDataPointsFilter"} <= set(shim.list_modules("DataPointsFilter"))
    assert {"CounterTransformationChecker", "DifferentialTransformationChecker", "BoundTransformationChecker"} <= set(shim.list_modules("TransformationChecker"))


def test_parameter_tables_match_the_cpu_twins(shim):
    """Same names / defaults / bounds as MatchersImpl.h:97-100, OutlierFiltersImpl.h:137, PointToPlane.h:74-75, VoxelGrid.h:69-73,
    SurfaceNormal.h:69-80, TransformationCheckersImpl.h (+ `device`, `cell`)."""
    p = shim.module_parameters("Matcher", "CudaKDTreeMatcher")
    assert p["knn"] == ("1", "1", "2147483647") and p["epsilon"] == ("0", "0", "inf") and p["searchType"] == ("1", "0", "2") and p["maxDist"] == ("inf", "0", "inf")
    assert shim.module_parameters("OutlierFilter", "CudaTrimmedDistOutlierFilter")["ratio"] == ("0.85", "0.0000001", "1.0")
    assert shim.module_parameters("OutlierFilter", "CudaSurfaceNormalOutlierFilter")["maxAngle"][0] == "1.57"
    assert shim.module_parameters("OutlierFilter", "CudaMaxDistOutlierFilter")["maxDist"] == ("1", "0.0000001", "inf")      # OutlierFiltersImpl.h:46-49
    assert shim.module_parameters("OutlierFilter", "CudaMinDistOutlierFilter")["minDist"] == ("1", "0.0000001", "inf")      # :64-67
    assert shim.module_parameters("OutlierFilter", "CudaMedianDistOutlierFilter")["factor"] == ("3", "0.0000001", "inf")    # :83-86
    p = shim.module_parameters("ErrorMinimizer", "CudaPointToPlaneErrorMinimizer")
    assert p["force2D"] == ("0", "0", "1") and p["force4DOF"] == ("0", "0", "1")
    p = shim.module_parameters("DataPointsFilter", "CudaVoxelGridDataPointsFilter")
    assert p["vSizeX"] == ("1.0", "0.001", "+inf") and p["useCentroid"][0] == "1" and p["averageExistingDescriptors"][0] == "1"
    p = shim.module_parameters("DataPointsFilter", "CudaSurfaceNormalDataPointsFilter")
    assert p["knn"] == ("5", "3", "2147483647") and p["keepNormals"][0] == "1" and p["keepSurfaceDescriptors"][0] == "1" and p["keepDensities"][0] == "0"
    p = shim.module_parameters("OutlierFilter", "CudaVarTrimmedDistOutlierFilter")                                           # OutlierFiltersImpl.h:153-159
    assert p["minRatio"] == ("0.05", "0.0000001", "1") and p["maxRatio"] == ("0.99", "0.0000001", "1") and p["lambda"][0] == "2.35"
    p = shim.module_parameters("ErrorMinimizer", "CudaPointToPlaneWithCovErrorMinimizer")                                   # PointToPlaneWithCov.h:70-76
    assert p["force2D"] == ("0", "0", "1") and p["sensorStdDev"] == ("0.01", "0.", "inf")
    assert set(shim.module_parameters("ErrorMinimizer", "CudaPointToPointErrorMinimizer")) == {"device"}                    # PointToPoint.h: no parameters
    assert shim.module_parameters("DataPointsFilter", "CudaMaxDistDataPointsFilter")["maxDist"] == ("1", "-inf", "inf")     # MaxDist.h:57-62
    p = shim.module_parameters("DataPointsFilter", "CudaMinDistDataPointsFilter")                                           # MinDist.h:57-62
    assert p["dim"] == ("-1", "-1", "2") and p["minDist"] == ("1", "-inf", "inf")
    p = shim.module_parameters("DataPointsFilter", "CudaBoundingBoxDataPointsFilter")                                       # BoundingBox.h:58-68
    assert p["xMin"] == ("-1", "-inf", "inf") and p["zMax"] == ("1", "-inf", "inf") and p["removeInside"] == ("1", "0", "1")
    p = shim.module_parameters("DataPointsFilter", "CudaRandomSamplingDataPointsFilter")                                    # RandomSampling.h:59-64
    assert p["prob"] == ("0.75", "0", "1") and p["randomSamplingMethod"] == ("0", "0", "1")
    p = shim.module_parameters("TransformationChecker", "DifferentialTransformationChecker")
    assert p["minDiffRotErr"][0] == "0.001" and p["smoothLength"][0] == "3"
    assert shim.module_parameters("TransformationChecker", "CounterTransformationChecker")["maxIterationCount"][0] == "40"


def test_open3d_slam_icp_yaml_loads_with_cuda_names(shim):
    icp = shim.ICP(to_cuda(yaml_text("open3d_slam_icp.yaml")))
    d = icp.describe()
    assert ("matcher", "CudaKDTreeMatcher") in d and ("errorMinimizer", "CudaPointToPlaneErrorMinimizer") in d
    assert [c for s, c in d if s == "outlierFilters"] == ["CudaTrimmedDistOutlierFilter", "CudaSurfaceNormalOutlierFilter"]
    assert [c for s, c in d if s == "transformationCheckers"] == ["DifferentialTransformationChecker", "CounterTransformationChecker"]
    assert not [c for s, c in d if s.endswith("DataPointsFilters")]      # both filter sections are empty in the shipped file


def test_reference_chain_files_parse(shim):
    # force4DOF golden config: its ObservationDirection / OrientNormals filters are out of scope -> drop those two entries
    t = to_cuda(yaml_text("force4DOFForPointToPlaneMinimizer.yaml"))
    t = t.replace("  - ObservationDirectionDataPointsFilter\n", "").replace("  - OrientNormalsDataPointsFilter:\n      towardCenter: 1\n", "")
    t = t.replace("      keepDensities: 1\n", "      keepDensities: 1\n      keepSurfaceDescriptors: 0\n")
    d = shim.ICP(t).describe()
    assert ("readingDataPointsFilters", "CudaSurfaceNormalDataPointsFilter") in d and ("referenceDataPointsFilters", "CudaSurfaceNormalDataPointsFilter") in d
    assert [c for s, c in d if s == "transformationCheckers"] == ["DifferentialTransformationChecker", "CounterTransformationChecker", "BoundTransformationChecker"]
    # icp_point_to_plane.yaml uses "{}" empty sections and MirrorMatcher (a test fixture, not registered here)
    with pytest.raises(shim.InvalidElement, match="MirrorMatcher"):
        shim.ICP(to_cuda(yaml_text("icp_point_to_plane.yaml")))
    ok = to_cuda(yaml_text("icp_point_to_plane.yaml")).replace("MirrorMatcher", "CudaKDTreeMatcher")
    assert ("matcher", "CudaKDTreeMatcher") in shim.ICP(ok).describe()


def test_bad_configs_raise_the_reference_exception_types(shim):
    """examples/data/unit_tests/badIcpConfig_*.yaml -> InvalidModuleType (ICP.cpp:156) / InvalidParameter (Registrar.h:89)."""
    def strip_unsupported(t):   # the two random samplers of the default chain are out of scope: remove their entries
        t = re.sub(r"  - RandomSamplingDataPointsFilter:\n(      .*\n)*", "", t)
        t = re.sub(r"  - SamplingSurfaceNormalDataPointsFilter:\n(      .*\n)*", "", t)
        return to_cuda(t)
    with pytest.raises(shim.InvalidModuleType, match="FAKE_MODULE_NAME"):
        shim.ICP(strip_unsupported(yaml_text("badIcpConfig_InvalidModuleType.yaml")).replace("FAKE_MODULE_NAME:\n", "FAKE_MODULE_NAME:\n  - IdentityDataPointsFilter\n"))
    bad = to_cuda(yaml_text("badIcpConfig_InvalidParameter.yaml"))
    bad = bad.replace("RandomSamplingDataPointsFilter", "CudaVoxelGridDataPointsFilter").replace("      prob: 0.5\n", "      vSizeX: 0.5\n")
    bad = re.sub(r"  - SamplingSurfaceNormalDataPointsFilter:\n(      .*\n)*", "", bad)
    with pytest.raises(shim.InvalidParameter, match="FAKE_PARAM"):
        shim.ICP(bad)
    # unknown module name -> InvalidElement (Registrar.h:146-158)
    with pytest.raises(shim.InvalidElement):
        shim.ICP("matcher:\n  NoSuchMatcher\n")
    # out-of-bounds parameter -> InvalidParameter with the reference's message (Parametrizable.cpp:183-186)
    with pytest.raises(shim.InvalidParameter, match="larger than maximum admissible value"):
        shim.ICP("outlierFilters:\n  - CudaTrimmedDistOutlierFilter:\n      ratio: 1.5\n")
    with pytest.raises(shim.InvalidParameter, match="smaller than minimum admissible value"):
        shim.ICP("referenceDataPointsFilters:\n  - CudaSurfaceNormalDataPointsFilter:\n      knn: 2\n")
    # "inf" is accepted for float parameters (Parametrizable.h:53-64)
    shim.ICP("matcher:\n  CudaKDTreeMatcher:\n    maxDist: inf\n    knn: 3\n")


def test_missing_modules_are_reported_before_running(shim):
    icp = shim.ICP("matcher:\n  CudaKDTreeMatcher\n")
    pts = np.zeros((4, 4), dtype=np.float32)
    with pytest.raises(shim.PmError, match="error minimizer"):      # ICP.cpp:267-268
        icp(pts, pts)


# ---------------------------------------------------------------------------------------------------------------
# GPU: chains loaded from YAML, module-by-module vs fast lane vs oracle
# ---------------------------------------------------------------------------------------------------------------
def _assert_same_T(Ta, Tb, tol=1e-5):
    assert np.abs(Ta[:3, 3].astype(np.float64) - Tb[:3, 3]).max() <= tol, (Ta, Tb)
    assert np.linalg.norm(Ta[:3, :3].astype(np.float64) - Tb[:3, :3]) / np.sqrt(2) <= tol, (Ta, Tb)


@pytest.mark.gpu
def test_open3d_slam_chain_generic_vs_fast_vs_oracle(shim, oracle, golden):
    car = golden["car"]
    ref, rn, q = oracle.feat4(car["ref_xyz"]), car["ref_normals"], oracle.feat4(car["data_xyz"])
    qn = oracle.surface_normals(q, 10, nthreads=8, orient_origin=[0, 0, 0])["normals"]
    icp = shim.ICP(to_cuda(yaml_text("open3d_slam_icp.yaml")))
    T_fast, s_fast = icp(q, ref, None, qn, rn)
    T_gen, s_gen = icp(q, ref, None, qn, rn, force_generic=True)
    assert s_fast["fast_lane"] and not s_gen["fast_lane"]
    assert s_fast["iterations"] == s_gen["iterations"]
    _assert_same_T(T_fast, T_gen, 1e-6)
    cfg = oracle.icp_config(knn=1, max_dist=0.5, epsilon=0.01, trim_ratio=0.9, max_angle=1.57, max_iter=15, min_diff_rot=0.001,
                            min_diff_trans=0.01, smooth_length=3, precision=oracle.ORC_MIXED, nthreads=8)
    code, T_o, st_o = oracle.icp(cfg, ref, rn, q, qn)
    assert code == 0 and st_o.iterations == s_fast["iterations"]
    _assert_same_T(T_fast, T_o)
    assert s_fast["overlap"] == pytest.approx(st_o.weighted_point_used_ratio, rel=1e-6)
    assert s_gen["overlap"] == pytest.approx(st_o.weighted_point_used_ratio, rel=1e-6)


@pytest.mark.gpu
def test_default_chain_on_car_clouds(shim, oracle, golden):
    """BASELINE config #1 through the registrable-module surface: setDefault() (ICP.cpp:96-109 with the CUDA twins)."""
    car = golden["car"]
    ref, rn, q = oracle.feat4(car["ref_xyz"]), car["ref_normals"], oracle.feat4(car["data_xyz"])
    icp = shim.ICP()
    icp.set_default()
    T, s = icp(q, ref, None, None, rn)
    # module defaults (TransformationCheckersImpl.h:90-92): minDiffRotErr 0.001, minDiffTransErr 0.001, smoothLength 3; Counter 40; ratio 0.85
    code, T_o, st_o = oracle.icp(oracle.icp_config(min_diff_trans=0.001, precision=oracle.ORC_MIXED, nthreads=8), ref, rn, q)
    assert s["fast_lane"] and s["iterations"] == st_o.iterations
    _assert_same_T(T, T_o)
    V = golden["T"]["validT3d"]
    assert abs(np.linalg.norm(V[:3, 3]) - np.linalg.norm(T[:3, 3])) < 0.1


@pytest.mark.gpu
def test_force4dof_chain_with_normal_filters(shim, oracle, golden):
    """The golden config that chains every in-scope module: SurfaceNormal filters on both clouds, KDTree knn 10,
    TrimmedDist 0.8 x SurfaceNormalOutlier 0.42, PointToPlane force4DOF, three checkers.  Its two out-of-scope filters
    (ObservationDirection / OrientNormals) are dropped, so the oracle is given the same un-oriented normals."""
    c0, c1 = golden["vtk"]["cloud0"], golden["vtk"]["cloud1"]
    t = to_cuda(yaml_text("force4DOFForPointToPlaneMinimizer.yaml"))
    t = t.replace("  - ObservationDirectionDataPointsFilter\n", "").replace("  - OrientNormalsDataPointsFilter:\n      towardCenter: 1\n", "")
    icp = shim.ICP(t)
    ref, q = oracle.feat4(c0), oracle.feat4(c1)
    T_fast, s_fast = icp(q, ref)
    T_gen, s_gen = icp(q, ref, force_generic=True)
    assert s_fast["fast_lane"] and not s_gen["fast_lane"] and s_fast["iterations"] == s_gen["iterations"]
    _assert_same_T(T_fast, T_gen, 1e-6)
    assert abs(T_fast[2, 2] - 1) < 1e-6 and np.abs(T_fast[2, :2]).max() < 1e-6 and np.abs(T_fast[:2, 2]).max() < 1e-6      # yaw only
    # oracle with normals taken from the CUDA filter (normal signs are unpinned; the gate depends on them)
    from open3d_slam_advanced_rss_2024_public_b200 import pm_shim
    _, rn, _ = pm_shim.apply_filter("CudaSurfaceNormalDataPointsFilter", {"knn": 15}, ref)
    _, qn, labels = pm_shim.apply_filter("CudaSurfaceNormalDataPointsFilter", {"knn": 12, "epsilon": 0.3, "keepNormals": 1, "keepDensities": 1}, q)
    assert [l[0] for l in labels] == ["normals", "densities", "linearity", "planarity", "curvature"]      # SurfaceNormal.cpp:126-144 order
    cfg = oracle.icp_config(knn=10, max_dist=1.5, epsilon=1.0, trim_ratio=0.8, max_angle=0.42, force4DOF=1, max_iter=100, min_diff_rot=0.001,
                            min_diff_trans=0.01, smooth_length=2, bound_rot=0.5, bound_trans=2.0, precision=oracle.ORC_MIXED, nthreads=8)
    code, T_o, st_o = oracle.icp(cfg, ref, rn, q, qn)
    assert code == 0 and st_o.iterations == s_fast["iterations"]
    _assert_same_T(T_fast, T_o)
    # the reference's golden transform for this config, at its own 5 % bar (GeneralTests.cpp:136-148)
    G = golden["T"]["force4DOFForPointToPlaneMinimizer"]
    P = np.concatenate([c1.astype(np.float64), np.ones((c1.shape[0], 1))], axis=1).T
    err = np.linalg.norm((T_fast.astype(np.float64) @ P - G @ P)[:3], axis=0)
    assert np.median(err) / np.median(np.linalg.norm((T_fast.astype(np.float64) @ P)[:3], axis=0)) < 0.05


@pytest.mark.gpu
def test_filters_by_name_match_oracle(shim, oracle, golden):
    """utest/ui/DataFilters.cpp shapes: a filter created through the registrar with string parameters."""
    pts = oracle.feat4(golden["car"]["data_xyz"])
    f, n, labels = shim.apply_filter("CudaVoxelGridDataPointsFilter", {"vSizeX": 0.2, "vSizeY": 0.2, "vSizeZ": 0.2}, pts)
    m, f_o, _, _ = oracle.voxelgrid(pts, None, 0.2, 0.2, 0.2)
    assert f.shape[0] == m and np.array_equal(f.view(np.uint32), f_o.view(np.uint32)) and labels == []
    f2, n2, labels = shim.apply_filter("CudaSurfaceNormalDataPointsFilter", {"knn": 5, "epsilon": 0.1, "keepNormals": 1, "keepDensities": 1,
                                                                           "keepEigenValues": 1}, pts)
    o = oracle.surface_normals(pts, 5, nthreads=8)
    assert [l[0] for l in labels] == ["normals", "densities", "eigValues", "linearity", "planarity", "curvature"]
    dot = np.abs(np.sum(n2.astype(np.float64) * o["normals"], axis=1))
    ok = ~np.all(o["normals"] == 0, axis=1)
    assert np.all(dot[ok] > 1 - 1e-6)      # fp32 storage of unit normals
    # the descriptors round 1 refused (SurfaceNormal.cpp:133-138): produced now, in the reference's label order
    f3, n3, labels = shim.apply_filter("CudaSurfaceNormalDataPointsFilter", {"knn": 5, "keepEigenVectors": 1, "keepMatchedIds": 1, "keepMeanDist": 1,
                                                                           "keepSurfaceDescriptors": 0, "smoothNormals": 1}, pts)
    assert labels == [("normals", 3), ("eigVectors", 9), ("matchedIds", 5), ("meanDists", 1)]
    sm = oracle.smooth_normals(pts, o["normals"], 5)
    assert np.abs(np.sum(n3.astype(np.float64) * sm, axis=1))[ok].min() > 0.0 and np.abs(n3 - sm)[ok].max() < 1e-4
    with pytest.raises(shim.InvalidParameter):
        shim.apply_filter("CudaSurfaceNormalDataPointsFilter", {"knn": 5, "keepNormals": 0, "smoothNormals": 1}, pts)


@pytest.mark.gpu
def test_chain_errors_surface_as_runtime_errors(shim, oracle, golden):
    """Mapper.cpp:420 catches std::runtime_error and keeps its initial guess: every failure must arrive as one."""
    car = golden["car"]
    ref, rn, q = oracle.feat4(car["ref_xyz"]), car["ref_normals"], oracle.feat4(car["data_xyz"])
    icp = shim.ICP(to_cuda(yaml_text("open3d_slam_icp.yaml")))
    far = np.eye(4, dtype=np.float32)
    far[:3, 3] = 500
    for generic in (False, True):
        with pytest.raises(shim.ConvergenceError):
            icp(q, ref, far, None, rn, force_generic=generic)
        with pytest.raises(shim.InvalidField):                      # reference without normals (PointToPlane.cpp:115)
            icp(q, ref, None, None, None, force_generic=generic)
    with pytest.raises(shim.PmError, match="empty"):                # ICP.cpp:357-359
        icp(np.zeros((0, 4), dtype=np.float32), ref, None, None, rn)


@pytest.mark.gpu
def test_chain_with_distance_outlier_filters(shim, oracle, golden):
    """A chain that adds the MaxDist / MedianDist twins runs in the generic lane (module by module).  With limits that reject
    nothing beyond what the trimmed filter already rejects, weights multiply to the same set and the result equals the plain chain;
    a tighter MaxDist than the matcher's radius shrinks the kept set."""
    car = golden["car"]
    ref, rn, q = oracle.feat4(car["ref_xyz"]), car["ref_normals"], oracle.feat4(car["data_xyz"])
    qn = oracle.surface_normals(q, 10, nthreads=8, orient_origin=[0, 0, 0])["normals"]
    base = to_cuda(yaml_text("open3d_slam_icp.yaml"))
    T0, s0 = shim.ICP(base)(q, ref, None, qn, rn, force_generic=True)
    loose = base.replace("outlierFilters:\n", "outlierFilters:\n  - CudaMaxDistOutlierFilter:\n      maxDist: 100.0\n  - CudaMedianDistOutlierFilter:\n      factor: 1000000.0\n", 1)
    assert loose != base
    T1, s1 = shim.ICP(loose)(q, ref, None, qn, rn)
    assert not s1["fast_lane"] and s1["iterations"] == s0["iterations"]
    _assert_same_T(T1, T0, 1e-6)
    tight = base.replace("outlierFilters:\n", "outlierFilters:\n  - CudaMaxDistOutlierFilter:\n      maxDist: 0.3\n", 1)
    T2, s2 = shim.ICP(tight)(q, ref, None, qn, rn)
    assert not s2["fast_lane"] and s2["overlap"] < s0["overlap"] and np.isfinite(T2).all()
    with pytest.raises(shim.InvalidParameter):
        shim.ICP(base.replace("outlierFilters:\n", "outlierFilters:\n  - CudaMedianDistOutlierFilter:\n      factor: 0\n", 1))


@pytest.mark.gpu
def test_compacting_filters_by_name_match_oracle(shim, oracle, golden):
    """utest/ui/DataFilters.cpp shapes (MaxDistDataPointsFilter, MinDistDataPointsFilter, BoundingBoxDataPointsFilter,
    RandomSamplingDataPointsFilter tests): created through the registrar with string parameters, membership against the oracle."""
    pts = oracle.feat4(golden["car"]["data_xyz"])
    cases = [("CudaMaxDistDataPointsFilter", {"dim": -1, "maxDist": 7.0}, dict(kind=0, dim=-1, dist=7.0)),
             ("CudaMaxDistDataPointsFilter", {"dim": 0, "maxDist": 0.5}, dict(kind=0, dim=0, dist=0.5)),
             ("CudaMinDistDataPointsFilter", {"dim": -1, "minDist": 5.0}, dict(kind=1, dim=-1, dist=5.0)),
             ("CudaMinDistDataPointsFilter", {"dim": 2, "minDist": 0.2}, dict(kind=1, dim=2, dist=0.2)),
             ("CudaBoundingBoxDataPointsFilter", {"xMin": -5.5, "xMax": 6.5, "yMin": -10, "yMax": 8, "zMin": -1, "zMax": 3, "removeInside": 1},
              dict(kind=2, box=(-5.5, 6.5, -10, 8, -1, 3), remove_inside=True)),
             ("CudaBoundingBoxDataPointsFilter", {"xMin": -5.5, "xMax": 6.5, "yMin": -10, "yMax": 8, "zMin": -1, "zMax": 3, "removeInside": 0},
              dict(kind=2, box=(-5.5, 6.5, -10, 8, -1, 3), remove_inside=False)),
             ("CudaRandomSamplingDataPointsFilter", {"prob": 0.5, "seed": 77}, dict(kind=3, prob=0.5, seed=77)),
             ("CudaRandomSamplingDataPointsFilter", {"prob": 0.25, "seed": 78, "randomSamplingMethod": 1}, dict(kind=3, prob=0.25, seed=78, random_method=1))]
    for name, params, okw in cases:
        f, _, _ = shim.apply_filter(name, params, pts)
        keep = oracle.cloud_filter(pts, **okw)
        assert 0 < keep.sum() < pts.shape[0], name
        assert np.array_equal(f[:, :3], pts[keep, :3]), (name, params)
    with pytest.raises(shim.InvalidParameter):
        shim.apply_filter("CudaMaxDistDataPointsFilter", {"dim": 3}, pts)                     # dim bound of the parameter table
    with pytest.raises(shim.InvalidParameter):
        shim.apply_filter("CudaRandomSamplingDataPointsFilter", {"prob": 1.5}, pts)


@pytest.mark.gpu
def test_chains_with_point_to_point_var_trimmed_and_covariance(shim, oracle, golden):
    """utest/ui/ErrorMinimizers.cpp / Outliers.cpp shapes (validate3dTransformation with the default chain and one module swapped):
    PointToPoint + VarTrimmedDist reach validT3d like the reference's own tests require; PointToPlaneWithCov returns the transform
    of the plain point-to-plane chain."""
    car = golden["car"]
    ref, rn, q = oracle.feat4(car["ref_xyz"]), car["ref_normals"], oracle.feat4(car["data_xyz"])
    base = """matcher:
  CudaKDTreeMatcher:
    knn: 1
outlierFilters:
  - OUTLIER
errorMinimizer:
  MINIMIZER
transformationCheckers:
  - CounterTransformationChecker:
      maxIterationCount: 40
  - DifferentialTransformationChecker:
      minDiffRotErr: 0.001
      minDiffTransErr: 0.01
      smoothLength: 3
inspector:
  NullInspector
logger:
  NullLogger
"""
    V = golden["T"]["validT3d"]
    trimmed = "CudaTrimmedDistOutlierFilter:\n      ratio: 0.85"
    vart = "CudaVarTrimmedDistOutlierFilter:\n      minRatio: 0.60\n      maxRatio: 0.80\n      lambda: 0.9"     # Outliers.cpp:113-123
    T_pp, s_pp = shim.ICP(base.replace("OUTLIER", trimmed).replace("MINIMIZER", "CudaPointToPointErrorMinimizer"))(q, ref, None, None, rn)
    T_vt, s_vt = shim.ICP(base.replace("OUTLIER", vart).replace("MINIMIZER", "CudaPointToPointErrorMinimizer"))(q, ref, None, None, rn)
    T_pl, s_pl = shim.ICP(base.replace("OUTLIER", trimmed).replace("MINIMIZER", "CudaPointToPlaneErrorMinimizer"))(q, ref, None, None, rn)
    T_cv, s_cv = shim.ICP(base.replace("OUTLIER", trimmed).replace("MINIMIZER", "CudaPointToPlaneWithCovErrorMinimizer:\n    sensorStdDev: 0.02"))(q, ref, None, None, rn)
    for T, s in ((T_pp, s_pp), (T_vt, s_vt), (T_pl, s_pl), (T_cv, s_cv)):
        assert np.isfinite(T).all() and not (s is not s_pl and s["fast_lane"])
        # validate3dTransformation (utest/utest.cpp): translation norm within 0.1 ... of validT3d
        assert abs(np.linalg.norm(V[:3, 3]) - np.linalg.norm(T[:3, 3])) < 0.1
        R = T[:3, :3].astype(np.float64) @ V[:3, :3].T
        assert np.degrees(np.arccos(np.clip((np.trace(R) - 1) / 2, -1, 1))) < 6.0
    assert s_cv["iterations"] == s_pl["iterations"]
    _assert_same_T(T_cv, T_pl, 1e-6)
    with pytest.raises(shim.InvalidParameter):
        shim.ICP(base.replace("OUTLIER", "CudaVarTrimmedDistOutlierFilter:\n      minRatio: 0.9\n      maxRatio: 0.5").replace("MINIMIZER", "CudaPointToPointErrorMinimizer"))
