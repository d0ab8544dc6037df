"""ctypes binding of oracle/liboracle.so — TEST INFRASTRUCTURE ONLY.

Imported only by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.
The product package never imports this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

ORC_REF32, ORC_MIXED = 0, 1
ORC_OK, ORC_E_CONVERGENCE, ORC_E_TRANSFORM, ORC_E_PARAM, ORC_E_EMPTY, ORC_E_FIELD = 0, -1, -2, -3, -4, -5
CROP_MAX_RADIUS, CROP_MIN_RADIUS, CROP_MINMAX_RADIUS, CROP_CYLINDER = 0, 1, 2, 3


class IcpConfig(C.Structure):
    _fields_ = [("matcher_kind", C.c_int32), ("knn", C.c_int32), ("max_dist", C.c_float), ("epsilon", C.c_float),
                ("trim_ratio", C.c_float), ("max_angle", C.c_float), ("force2D", C.c_int32), ("force4DOF", C.c_int32),
                ("max_iter", C.c_int32), ("min_diff_rot", C.c_float), ("min_diff_trans", C.c_float),
                ("smooth_length", C.c_int32), ("bound_rot", C.c_float), ("bound_trans", C.c_float),
                ("precision", C.c_int32), ("nthreads", C.c_int32)]


class IcpStats(C.Structure):
    _fields_ = [("iterations", C.c_int32), ("max_iterations_reached", C.c_int32), ("status", C.c_int32),
                ("kept_last", C.c_int32), ("points_touched", C.c_uint64), ("point_used_ratio", C.c_float),
                ("weighted_point_used_ratio", C.c_float), ("last_tau", C.c_float), ("mean_ref", C.c_float * 3),
                ("t_reference_s", C.c_double), ("t_converge_s", C.c_double)]


class CropSpec(C.Structure):
    _fields_ = [("kind", C.c_int32), ("invert", C.c_int32), ("center", C.c_double * 3), ("rmin", C.c_double),
                ("rmax", C.c_double), ("zmin", C.c_double), ("zmax", C.c_double)]


def build(force=False):
    so = os.path.join(_HERE, "liboracle.so")
    src = os.path.join(_HERE, "oracle.cpp")
    if force or not os.path.exists(so) or (os.path.exists(src) and os.path.getmtime(src) > os.path.getmtime(so)):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return so


def lib():
    global _LIB
    if _LIB is None:
        os.environ.setdefault("OMP_WAIT_POLICY", "passive")   # spinning workers starve each other on shared hosts
        _LIB = C.CDLL(build())
        L = _LIB
        L.orc_kdtree_build.restype = C.c_void_p
        L.orc_kdtree_build.argtypes = [C.c_void_p, C.c_int32, C.c_int64]
        L.orc_kdtree_free.argtypes = [C.c_void_p]
        L.orc_kdtree_knn.restype = C.c_uint64
        L.orc_kdtree_knn.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_int64, C.c_int32, C.c_float, C.c_int32,
                                     C.c_void_p, C.c_void_p, C.c_int32]
        L.orc_knn_brute.restype = C.c_uint64
        L.orc_knn_brute.argtypes = [C.c_void_p, C.c_int32, C.c_int64, C.c_void_p, C.c_int32, C.c_int64, C.c_int32, C.c_float,
                                    C.c_void_p, C.c_void_p, C.c_int32]
        L.orc_transform.restype = C.c_int32
        L.orc_transform.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_int64]
        L.orc_trimmed_dist.restype = C.c_int32
        L.orc_trimmed_dist.argtypes = [C.c_void_p, C.c_int64, C.c_float, C.c_void_p, C.c_void_p]
        L.orc_dist_outlier.argtypes = [C.c_int32, C.c_void_p, C.c_int64, C.c_float, C.c_void_p, C.c_void_p]
        L.orc_surface_normal_outlier.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int64, C.c_float, C.c_void_p]
        L.orc_no_filter_weights.argtypes = [C.c_void_p, C.c_int64, C.c_void_p]
        L.orc_point_to_plane.restype = C.c_int32
        L.orc_point_to_plane.argtypes = [C.c_void_p, C.c_int32, C.c_int64, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p,
                                         C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_void_p,
                                         C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_icp.restype = C.c_int32
        L.orc_icp.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_int64, C.c_void_p, C.c_int32, C.c_void_p,
                              C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_icp_ref_create.restype = C.c_void_p
        L.orc_icp_ref_create.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_int64]
        L.orc_icp_ref_free.argtypes = [C.c_void_p]
        L.orc_icp_compute.restype = C.c_int32
        L.orc_icp_compute.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_int64, C.c_void_p,
                                      C.c_void_p, C.c_void_p]
        L.orc_voxelgrid.restype = C.c_int64
        L.orc_voxelgrid.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_int32, C.c_int64, C.c_float, C.c_float, C.c_float,
                                    C.c_int32, C.c_int32, C.c_void_p]
        L.orc_surface_normals.restype = C.c_int32
        L.orc_surface_normals.argtypes = [C.c_void_p, C.c_int32, C.c_int64, C.c_int32, C.c_float, C.c_int32, C.c_int32,
                                          C.c_void_p] + [C.c_void_p] * 8
        L.orc_crop.restype = C.c_int64
        L.orc_crop.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_transform_f64.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]
        L.orc_voxelize_within_volume.restype = C.c_int64
        L.orc_voxelize_within_volume.argtypes = [C.c_double, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p,
                                                 C.c_void_p, C.c_void_p]
        L.orc_voxel_down_sample.restype = C.c_int64
        L.orc_voxel_down_sample.argtypes = [C.c_double, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]
        L.orc_o3d_to_pm.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]
    return _LIB


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f32(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float32)


def _f64(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float64)


def feat4(xyz):
    """N x 3 -> N x 4 float32 (x y z 1): the memory layout of Eigen's column-major 4 x N `features`."""
    xyz = np.asarray(xyz, dtype=np.float32)
    out = np.ones((xyz.shape[0], 4), dtype=np.float32)
    out[:, :3] = xyz[:, :3]
    return out


class KdTree:
    def __init__(self, feat):
        self.feat = _f32(feat)
        self.n, self.stride = self.feat.shape
        self.h = lib().orc_kdtree_build(_p(self.feat), self.stride, self.n)

    def knn(self, q, k=1, max_dist=np.inf, allow_self=True, nthreads=1):
        q = _f32(q)
        ids = np.empty((q.shape[0], k), dtype=np.int32)
        d2 = np.empty((q.shape[0], k), dtype=np.float32)
        visits = lib().orc_kdtree_knn(self.h, _p(q), q.shape[1], q.shape[0], k, max_dist, int(allow_self), _p(ids), _p(d2), nthreads)
        return ids, d2, visits

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_kdtree_free(self.h)
            self.h = None


def knn_brute(ref, q, k=1, max_dist=np.inf, nthreads=8):
    ref, q = _f32(ref), _f32(q)
    ids = np.empty((q.shape[0], k), dtype=np.int32)
    d2 = np.empty((q.shape[0], k), dtype=np.float32)
    lib().orc_knn_brute(_p(ref), ref.shape[1], ref.shape[0], _p(q), q.shape[1], q.shape[0], k, max_dist, _p(ids), _p(d2), nthreads)
    return ids, d2


def transform(T, feat, normals=None):
    T = _f32(T).reshape(16)
    feat = _f32(feat).copy()
    normals = None if normals is None else _f32(normals).copy()
    st = lib().orc_transform(_p(T), _p(feat), feat.shape[1], _p(normals), feat.shape[0])
    return st, feat, normals


def trimmed_dist(d2, ratio):
    d2 = _f32(d2)
    tau = C.c_float(0)
    w = np.empty(d2.shape, dtype=np.float32)
    st = lib().orc_trimmed_dist(_p(d2), d2.size, ratio, C.byref(tau), _p(w))
    return st, tau.value, w


def dist_outlier(kind, d2, param):
    """MaxDist (0) / MinDist (1) / MedianDist (2) outlier filters: returns (status, limit, weights)."""
    d2 = _f32(d2)
    lim = C.c_float(0)
    w = np.empty(d2.shape, dtype=np.float32)
    st = lib().orc_dist_outlier(int(kind), _p(d2), d2.size, float(param), C.byref(lim), _p(w))
    return st, lim.value, w


def surface_normal_outlier(nrm_read, nrm_ref, ids, max_angle):
    ids = np.ascontiguousarray(ids, dtype=np.int32)
    n, k = ids.shape
    w = np.empty((n, k), dtype=np.float32)
    lib().orc_surface_normal_outlier(_p(_f32(nrm_read)), _p(_f32(nrm_ref)), _p(ids), k, n, max_angle, _p(w))
    return w


def no_filter_weights(d2):
    d2 = _f32(d2)
    w = np.empty(d2.shape, dtype=np.float32)
    lib().orc_no_filter_weights(_p(d2), d2.size, _p(w))
    return w


def point_to_plane(read_feat, ref_feat, ref_normals, ids, d2, w, force2D=False, force4DOF=False, precision=ORC_MIXED):
    read_feat, ref_feat, ref_normals = _f32(read_feat), _f32(ref_feat), _f32(ref_normals)
    ids = np.ascontiguousarray(ids, dtype=np.int32)
    d2, w = _f32(d2), _f32(w)
    n = read_feat.shape[0]
    k = ids.size // n
    dim = 3 if force2D else (4 if force4DOF else 6)
    T = np.empty(16, dtype=np.float32)
    A = np.zeros(dim * dim, dtype=np.float64)
    b = np.zeros(dim, dtype=np.float64)
    x = np.zeros(dim, dtype=np.float64)
    kept = C.c_int32(0)
    pur, wpur = C.c_float(0), C.c_float(0)
    st = lib().orc_point_to_plane(_p(read_feat), read_feat.shape[1], n, _p(ref_feat), ref_feat.shape[1], _p(ref_normals), _p(ids),
                                  _p(d2), _p(w), k, int(force2D), int(force4DOF), precision, _p(T), _p(A), _p(b), _p(x),
                                  C.byref(kept), C.byref(pur), C.byref(wpur))
    return dict(status=st, T=T.reshape(4, 4), A=A.reshape(dim, dim), b=b, x=x, kept=kept.value,
                point_used_ratio=pur.value, weighted_point_used_ratio=wpur.value)


def icp_config(**kw):
    c = IcpConfig(matcher_kind=0, knn=1, max_dist=float("inf"), epsilon=0.0, trim_ratio=0.85, max_angle=-1.0, force2D=0,
                  force4DOF=0, max_iter=40, min_diff_rot=0.001, min_diff_trans=0.01, smooth_length=3, bound_rot=-1.0,
                  bound_trans=-1.0, precision=ORC_REF32, nthreads=1)
    for k, v in kw.items():
        if not hasattr(c, k):
            raise AttributeError(k)
        setattr(c, k, v)
    return c


def icp(cfg, ref_feat, ref_normals, read_feat, read_normals=None, T_init=None):
    ref_feat, ref_normals, read_feat, read_normals = _f32(ref_feat), _f32(ref_normals), _f32(read_feat), _f32(read_normals)
    T_init = np.eye(4, dtype=np.float32) if T_init is None else _f32(T_init)
    T = np.zeros(16, dtype=np.float32)
    st = IcpStats()
    code = lib().orc_icp(C.byref(cfg), _p(ref_feat), ref_feat.shape[1], _p(ref_normals), ref_feat.shape[0], _p(read_feat),
                         read_feat.shape[1], _p(read_normals), read_feat.shape[0], _p(T_init.reshape(16)), _p(T), C.byref(st))
    return code, T.reshape(4, 4), st


class IcpReference:
    """ICP::initReference done once (Mapper.cpp:363), then compute() many times (Mapper.cpp:391-393)."""

    def __init__(self, cfg, ref_feat, ref_normals):
        self.cfg = cfg
        ref_feat, ref_normals = _f32(ref_feat), _f32(ref_normals)
        self.h = lib().orc_icp_ref_create(C.byref(cfg), _p(ref_feat), ref_feat.shape[1], _p(ref_normals), ref_feat.shape[0])

    def compute(self, read_feat, read_normals=None, T_init=None, cfg=None):
        cfg = self.cfg if cfg is None else cfg
        read_feat, read_normals = _f32(read_feat), _f32(read_normals)
        T_init = np.eye(4, dtype=np.float32) if T_init is None else _f32(T_init)
        T = np.zeros(16, dtype=np.float32)
        st = IcpStats()
        code = lib().orc_icp_compute(C.byref(cfg), self.h, _p(read_feat), read_feat.shape[1], _p(read_normals),
                                     read_feat.shape[0], _p(T_init.reshape(16)), _p(T), C.byref(st))
        return code, T.reshape(4, 4), st

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_icp_ref_free(self.h)
            self.h = None


def voxelgrid(feat, desc, vx, vy, vz, use_centroid=True, average_descriptors=True):
    feat = _f32(feat).copy()
    desc = None if desc is None else _f32(desc).copy()
    n = feat.shape[0]
    cells = np.empty(n, dtype=np.uint32)
    m = lib().orc_voxelgrid(_p(feat), feat.shape[1], _p(desc), 0 if desc is None else desc.shape[1], n, vx, vy, vz,
                            int(use_centroid), int(average_descriptors), _p(cells))
    if m < 0:
        return m, None, None, cells
    return m, feat[:m], (None if desc is None else desc[:m]), cells


def surface_normals(feat, knn, max_dist=np.inf, precision=ORC_MIXED, nthreads=8, orient_origin=None, keep=("normals",)):
    feat = _f32(feat)
    n = feat.shape[0]
    shapes = dict(normals=(n, 3), densities=(n,), eig_values=(n, 3), eig_vectors=(n, 9), matched_ids=(n, knn),
                  mean_dists=(n,), lin_plan_curv=(n, 3))
    out = {k: (np.zeros(s, dtype=np.float32) if k in keep else None) for k, s in shapes.items()}
    deg = C.c_int32(0)
    oo = None if orient_origin is None else _f32(orient_origin)
    st = lib().orc_surface_normals(_p(feat), feat.shape[1], n, knn, max_dist, precision, nthreads, _p(oo), _p(out["normals"]),
                                   _p(out["densities"]), _p(out["eig_values"]), _p(out["eig_vectors"]), _p(out["matched_ids"]),
                                   _p(out["mean_dists"]), _p(out["lin_plan_curv"]), C.byref(deg))
    out["status"] = st
    out["degenerate"] = deg.value
    return out


def point_to_plane_residual(read_feat, ref_feat, ref_normals, ids, d2, w, force2D=False, precision=ORC_MIXED):
    """PointToPlaneErrorMinimizer::getResidualError: sum of w ((reading - reference) . n)^2 over the kept pairs."""
    read_feat, ref_feat, ref_normals = _f32(read_feat), _f32(ref_feat), _f32(ref_normals)
    ids = np.ascontiguousarray(ids, dtype=np.int32)
    d2, w = _f32(d2), _f32(w)
    k = ids.size // max(read_feat.shape[0], 1)
    out = C.c_double(0)
    L = lib()
    L.orc_point_to_plane_residual.restype = C.c_int32
    L.orc_point_to_plane_residual.argtypes = [C.c_void_p, C.c_int32, C.c_int64, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                              C.c_int32, C.c_int32, C.c_int32, C.POINTER(C.c_double)]
    st = L.orc_point_to_plane_residual(_p(read_feat), read_feat.shape[1], read_feat.shape[0], _p(ref_feat), ref_feat.shape[1], _p(ref_normals),
                                       _p(ids), _p(d2), _p(w), k, int(force2D), precision, C.byref(out))
    assert st == 0, st
    return out.value


def estimate_normals_f64(xyz, knn, radius, origin=None, exact=None, nthreads=8):
    """Open3D EstimateNormals(Hybrid) + OrientNormalsTowardsCameraLocation in fp64; exact=None: brute-force neighbours up to 20 000 points."""
    xyz = _f64(xyz)
    n = xyz.shape[0]
    exact = (n <= 20_000) if exact is None else exact
    out = np.zeros((n, 3), dtype=np.float64)
    deg = C.c_int32(0)
    oo = None if origin is None else _f64(origin)
    L = lib()
    L.orc_estimate_normals_f64.restype = C.c_int32
    L.orc_estimate_normals_f64.argtypes = [C.c_void_p, C.c_int64, C.c_int32, C.c_double, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.POINTER(C.c_int32)]
    st = L.orc_estimate_normals_f64(_p(xyz), n, int(knn), float(radius), _p(oo), int(bool(exact)), int(nthreads), _p(out), C.byref(deg))
    assert st == 0, st
    return out, deg.value


def smooth_normals(feat, normals, knn, max_dist=np.inf, nthreads=8):
    """smoothNormals of SurfaceNormal.cpp:299-323 (sequential in-place sweep); returns the smoothed copy."""
    feat = _f32(feat)
    out = np.array(normals, dtype=np.float32, order="C", copy=True)
    L = lib()
    L.orc_smooth_normals.restype = C.c_int32
    L.orc_smooth_normals.argtypes = [C.c_void_p, C.c_int32, C.c_int64, C.c_int32, C.c_float, C.c_int32, C.c_void_p]
    st = L.orc_smooth_normals(_p(feat), feat.shape[1], feat.shape[0], knn, max_dist, nthreads, _p(out))
    assert st == 0, st
    return out


def crop_spec(kind, center, rmin=0.0, rmax=0.0, zmin=0.0, zmax=0.0, invert=False):
    s = CropSpec(kind=kind, invert=int(invert), rmin=rmin, rmax=rmax, zmin=zmin, zmax=zmax)
    for i in range(3):
        s.center[i] = float(center[i])
    return s


def crop(spec, xyz, nrm=None):
    xyz, nrm = _f64(xyz), _f64(nrm)
    n = xyz.shape[0]
    xo = np.empty_like(xyz)
    no = None if nrm is None else np.empty_like(nrm)
    mask = np.empty(n, dtype=np.uint8)
    m = lib().orc_crop(C.byref(spec), _p(xyz), _p(nrm), n, _p(xo), _p(no), _p(mask))
    return xo[:m], (None if no is None else no[:m]), mask.astype(bool)


def transform_f64(T, xyz, nrm=None):
    T, xyz, nrm = _f64(T).reshape(16), _f64(xyz), _f64(nrm)
    xo = np.empty_like(xyz)
    no = None if nrm is None else np.empty_like(nrm)
    lib().orc_transform_f64(_p(T), _p(xyz), _p(nrm), xyz.shape[0], _p(xo), _p(no))
    return xo, no


def voxelize_within_volume(voxel, spec, xyz, nrm=None):
    xyz, nrm = _f64(xyz), _f64(nrm)
    xo = np.empty_like(xyz)
    no = None if nrm is None else np.empty_like(nrm)
    npass = C.c_int64(0)
    m = lib().orc_voxelize_within_volume(voxel, C.byref(spec), _p(xyz), _p(nrm), xyz.shape[0], _p(xo), _p(no), C.byref(npass))
    return xo[:m], (None if no is None else no[:m]), npass.value


def carve(scan_xyz, sensor, map_xyz, map_nrm, spec, voxel=0.1, max_ray_length=20.0, truncation=0.1, min_dot=0.5):
    """Space carving (Submap.cpp:130-144, helpers.cpp:245-281): boolean mask of the map points to remove; scan in the map frame."""
    scan_xyz, map_xyz, map_nrm, sensor = _f64(scan_xyz), _f64(map_xyz), _f64(map_nrm), _f64(sensor)
    mask = np.zeros(map_xyz.shape[0], dtype=np.uint8)
    lib().orc_carve.restype = C.c_int64
    lib().orc_carve.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_double, C.c_double,
                                C.c_double, C.c_double, C.c_void_p]
    lib().orc_carve(_p(scan_xyz), scan_xyz.shape[0], _p(sensor), _p(map_xyz), _p(map_nrm), map_xyz.shape[0], C.byref(spec), voxel,
                    max_ray_length, truncation, min_dot, _p(mask))
    return mask.astype(bool)


def pointcloud2(data, n, point_step, fields):
    """PointCloud2 bytes -> (feat4 f32, normals f32 | None, xyz f64, normals f64 | None); fields = {name: byte offset}."""
    data = np.ascontiguousarray(np.frombuffer(data, dtype=np.uint8) if not isinstance(data, np.ndarray) else data.view(np.uint8).reshape(-1))
    offs = np.array([fields.get(k, -1) for k in ("x", "y", "z", "normal_x", "normal_y", "normal_z")], dtype=np.int32)
    has_n = bool((offs[3:] >= 0).all())
    f4, x64 = np.zeros((n, 4), dtype=np.float32), np.zeros((n, 3), dtype=np.float64)
    n3 = np.zeros((n, 3), dtype=np.float32) if has_n else None
    n64 = np.zeros((n, 3), dtype=np.float64) if has_n else None
    lib().orc_pointcloud2.argtypes = [C.c_void_p, C.c_int64, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    st = lib().orc_pointcloud2(_p(data), n, point_step, _p(offs), _p(f4), _p(n3), _p(x64), _p(n64))
    return st, f4, n3, x64, n64


def overlap(src, tgt, T, voxel, min_pts=1):
    """computeIndicesOfOverlappingPoints: ascending indices (source, target)."""
    src, tgt, T = _f64(src), _f64(tgt), _f64(T).reshape(16)
    sm, tm = np.zeros(src.shape[0], dtype=np.uint8), np.zeros(tgt.shape[0], dtype=np.uint8)
    lib().orc_overlap.restype = None
    lib().orc_overlap.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_void_p, C.c_double, C.c_int64, C.c_void_p, C.c_void_p]
    lib().orc_overlap(_p(src), src.shape[0], _p(tgt), tgt.shape[0], _p(T), voxel, min_pts, _p(sm), _p(tm))
    return np.nonzero(sm)[0], np.nonzero(tm)[0]


def information_matrix(src, tgt, max_dist, T):
    """GetInformationMatrixFromPointClouds (fp64 brute-force correspondences): (6 x 6, count)."""
    src, tgt, T = _f64(src), _f64(tgt), _f64(T).reshape(16)
    info = np.zeros(36, dtype=np.float64)
    lib().orc_information_matrix.restype = C.c_int64
    lib().orc_information_matrix.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_double, C.c_void_p, C.c_void_p]
    cnt = lib().orc_information_matrix(_p(src), src.shape[0], _p(tgt), tgt.shape[0], max_dist, _p(T), _p(info))
    return info.reshape(6, 6), int(cnt)


def voxel_down_sample(voxel, xyz, nrm=None):
    xyz, nrm = _f64(xyz), _f64(nrm)
    xo = np.empty_like(xyz)
    no = None if nrm is None else np.empty_like(nrm)
    m = lib().orc_voxel_down_sample(voxel, _p(xyz), _p(nrm), xyz.shape[0], _p(xo), _p(no))
    return xo[:m], (None if no is None else no[:m])


def o3d_to_pm(xyz, nrm=None):
    xyz, nrm = _f64(xyz), _f64(nrm)
    n = xyz.shape[0]
    f = np.empty((n, 4), dtype=np.float32)
    nn = None if nrm is None else np.empty((n, 3), dtype=np.float32)
    lib().orc_o3d_to_pm(_p(xyz), _p(nrm), n, _p(f), _p(nn))
    return f, nn


def var_trimmed_dist(d2, min_ratio, max_ratio, lam):
    """VarTrimmedDistOutlierFilter: returns (status, optimised ratio, limit, weights)."""
    d2 = _f32(d2)
    ratio, tau = C.c_float(0), C.c_float(0)
    w = np.empty(d2.shape, dtype=np.float32)
    L = lib()
    L.orc_var_trimmed_dist.argtypes = [C.c_void_p, C.c_int64, C.c_float, C.c_float, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p]
    st = L.orc_var_trimmed_dist(_p(d2), d2.size, min_ratio, max_ratio, lam, C.byref(ratio), C.byref(tau), _p(w))
    return st, ratio.value, tau.value, w


def point_to_point(read_feat, ref_feat, ids, d2, w, precision=ORC_MIXED):
    """PointToPointErrorMinimizer::compute on explicit matches / weights: returns (status, T 4x4 float32, kept)."""
    read_feat, ref_feat = _f32(read_feat), _f32(ref_feat)
    ids = np.ascontiguousarray(ids, dtype=np.int32)
    d2, w = _f32(d2), _f32(w)
    n, k = ids.shape
    T = np.zeros((4, 4), dtype=np.float32)
    kept = C.c_int32(0)
    L = lib()
    L.orc_point_to_point.argtypes = [C.c_void_p, C.c_int32, C.c_int64, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32,
                                     C.c_int32, C.c_void_p, C.c_void_p]
    st = L.orc_point_to_point(_p(read_feat), read_feat.shape[1], n, _p(ref_feat), ref_feat.shape[1], _p(ids), _p(d2), _p(w), k, precision,
                              _p(T), C.byref(kept))
    return st, T, kept.value


def point_to_plane_covariance(read_feat, ref_feat, ref_normals, ids, d2, w, T, sensor_std_dev, precision=ORC_MIXED):
    """PointToPlaneWithCovErrorMinimizer::estimateCovariance: returns (status, 6x6 float64)."""
    read_feat, ref_feat, ref_normals = _f32(read_feat), _f32(ref_feat), _f32(ref_normals)
    ids = np.ascontiguousarray(ids, dtype=np.int32)
    d2, w, T = _f32(d2), _f32(w), _f32(T)
    n, k = ids.shape
    cov = np.zeros((6, 6), dtype=np.float64)
    L = lib()
    L.orc_point_to_plane_covariance.argtypes = [C.c_void_p, C.c_int32, C.c_int64, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p,
                                                C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_float, C.c_void_p]
    st = L.orc_point_to_plane_covariance(_p(read_feat), read_feat.shape[1], n, _p(ref_feat), ref_feat.shape[1], _p(ref_normals), _p(ids),
                                         _p(d2), _p(w), k, precision, _p(T), sensor_std_dev, _p(cov))
    return st, cov


FILTER_MAX_DIST, FILTER_MIN_DIST, FILTER_BOUNDING_BOX, FILTER_RANDOM_SAMPLING = 0, 1, 2, 3


def cloud_filter(feat, kind, dim=-1, dist=1.0, box=(-1, 1, -1, 1, -1, 1), remove_inside=True, random_method=0, prob=0.75, seed=1):
    """MaxDist / MinDist / BoundingBox / RandomSampling DataPointsFilters: returns the boolean keep mask."""
    feat = _f32(feat)
    n = feat.shape[0]
    keep = np.zeros(n, dtype=np.uint8)
    b = np.asarray(box, dtype=np.float32)
    L = lib()
    L.orc_cloud_filter.restype = C.c_int64
    L.orc_cloud_filter.argtypes = [C.c_void_p, C.c_int32, C.c_int64, C.c_int32, C.c_int32, C.c_float, C.c_void_p, C.c_int32, C.c_int32,
                                   C.c_double, C.c_uint64, C.c_void_p]
    cnt = L.orc_cloud_filter(_p(feat), feat.shape[1], n, kind, dim, dist, _p(b), int(remove_inside), random_method, prob, seed, _p(keep))
    assert cnt == int(keep.sum())
    return keep.astype(bool)
