"""ctypes binding of libb200reg.so (include/b200reg.h) — the call a Python user makes.

There is no CPU fallback: if the CUDA library is missing or no GPU is visible, every entry point raises.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libb200reg.so")
_LIB = None

OK = 0
E_INVALID_PARAM, E_INVALID_FIELD, E_NO_POINTS, E_NOT_CONVERGED, E_TRANSFORM, E_CUDA, E_OOM = -1, -2, -3, -4, -5, -6, -7
CROP_MAX_RADIUS, CROP_MIN_RADIUS, CROP_MINMAX_RADIUS, CROP_CYLINDER = 0, 1, 2, 3


# exception types mirror libpointmatcher's (PointMatcher.h:86-103,156-159,304-307); all derive from RuntimeError
class B200RegError(RuntimeError):
    code = None


class InvalidParameter(B200RegError):
    code = E_INVALID_PARAM


class InvalidField(B200RegError):
    code = E_INVALID_FIELD


class NoPoints(B200RegError):
    code = E_NO_POINTS


class ConvergenceError(B200RegError):
    code = E_NOT_CONVERGED


class TransformationError(B200RegError):
    code = E_TRANSFORM


class CudaError(B200RegError):
    code = E_CUDA


class OutOfMemory(B200RegError):
    code = E_OOM


_EXC = {c.code: c for c in (InvalidParameter, InvalidField, NoPoints, ConvergenceError, TransformationError, CudaError, OutOfMemory)}


class StepStats(C.Structure):
    _fields_ = [("kept", C.c_int64), ("n_finite", C.c_int64), ("weight_sum", C.c_double), ("tau", C.c_float),
                ("point_used_ratio", C.c_float), ("weighted_point_used_ratio", C.c_float), ("visit_count", C.c_uint64),
                ("mean_reading", C.c_double * 3), ("mean_reference", C.c_double * 3)]


class CarveParams(C.Structure):
    """SpaceCarvingParameters (Parameters.hpp:88-95)"""
    _fields_ = [("voxel_size", C.c_double), ("max_raytracing_length", C.c_double), ("truncation_distance", C.c_double),
                ("min_dot_product_with_normal", C.c_double)]


class FilterParams(C.Structure):
    """b200reg_filter_params: MaxDist / MinDist / BoundingBox / RandomSampling DataPointsFilters"""
    _fields_ = [("kind", C.c_int32), ("dim", C.c_int32), ("dist", C.c_float), ("box", C.c_float * 6), ("remove_inside", C.c_int32),
                ("random_method", C.c_int32), ("prob", C.c_double), ("seed", C.c_uint64)]


FILTER_MAX_DIST, FILTER_MIN_DIST, FILTER_BOUNDING_BOX, FILTER_RANDOM_SAMPLING = 0, 1, 2, 3


class Pc2Layout(C.Structure):
    """The parts of sensor_msgs/PointCloud2 the deserialisers read (byte offsets of FLOAT32 fields; < 0 = absent)."""
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("point_step", C.c_uint32), ("off_x", C.c_int32), ("off_y", C.c_int32),
                ("off_z", C.c_int32), ("off_normal_x", C.c_int32), ("off_normal_y", C.c_int32), ("off_normal_z", C.c_int32)]


def pc2_layout(width, height, point_step, fields):
    """fields: {name: byte offset} as in sensor_msgs/PointField."""
    g = lambda k: int(fields.get(k, -1))  # noqa: E731
    return Pc2Layout(width, height, point_step, g("x"), g("y"), g("z"), g("normal_x"), g("normal_y"), g("normal_z"))


class IcpConfig(C.Structure):
    _fields_ = [("knn", C.c_int32), ("max_dist", C.c_float), ("epsilon", C.c_float), ("trim_ratio", C.c_float),
                ("max_angle", C.c_float), ("force2D", C.c_int32), ("force4DOF", C.c_int32), ("max_iter", C.c_int32),
                ("min_diff_rot", C.c_float), ("min_diff_trans", C.c_float), ("smooth_length", C.c_int32),
                ("bound_rot", C.c_float), ("bound_trans", C.c_float), ("cell", C.c_float), ("checker_order", C.c_int32 * 3)]


CHECKER_COUNTER, CHECKER_DIFFERENTIAL, CHECKER_BOUND = 1, 2, 3


class NormalsOptions(C.Structure):
    """b200reg_normals_options"""
    _fields_ = [("knn", C.c_int32), ("max_dist", C.c_float), ("epsilon", C.c_float), ("smooth_normals", C.c_int32), ("orient_origin", C.c_void_p),
                ("densities", C.c_void_p), ("eig_values", C.c_void_p), ("eig_vectors", C.c_void_p), ("matched_ids", C.c_void_p),
                ("mean_dists", C.c_void_p), ("lin_plan_curv", C.c_void_p)]


class HostCloud(C.Structure):
    """b200reg_host_cloud: a cloud in host memory (layout of b200reg_cloud_upload_f32)."""
    _fields_ = [("feat", C.c_void_p), ("feat_stride", C.c_int32), ("normals", C.c_void_p), ("n", C.c_int64)]


class IcpStats(C.Structure):
    _fields_ = [("iterations", C.c_int32), ("max_iterations_reached", C.c_int32), ("reference_point_count", C.c_int64),
                ("reading_point_count", C.c_int64), ("points_touched", C.c_uint64), ("overlap_ratio", C.c_float),
                ("point_used_ratio", C.c_float), ("last_tau", C.c_float), ("kept_last", C.c_int64),
                ("reference_preprocessing_s", C.c_double), ("reading_preprocessing_s", C.c_double),
                ("convergence_s", C.c_double), ("device_ms", C.c_double)]


class CropSpec(C.Structure):
    _fields_ = [("kind", C.c_int32), ("invert", C.c_int32), ("center", C.c_double * 3), ("rmin", C.c_double),
                ("rmax", C.c_double), ("zmin", C.c_double), ("zmax", C.c_double)]


# every symbol include/b200reg.h declares: name -> (restype, argtypes)
_VP = C.c_void_p
SYMBOLS = {
    "b200reg_abi_version": (C.c_int32, []),
    "b200reg_ctx_create": (C.c_int32, [C.c_int32, C.POINTER(_VP)]),
    "b200reg_ctx_destroy": (None, [_VP]),
    "b200reg_last_error": (C.c_char_p, [_VP]),
    "b200reg_ctx_sync": (C.c_int32, [_VP]),
    "b200reg_ctx_device": (C.c_int32, [_VP]),
    "b200reg_ctx_launch_count": (C.c_uint64, [_VP]),
    "b200reg_ctx_timer_start": (C.c_int32, [_VP]),
    "b200reg_ctx_timer_stop": (C.c_int32, [_VP, C.POINTER(C.c_float)]),
    "b200reg_ctx_flush_l2": (C.c_int32, [_VP]),
    "b200reg_ctx_set_profiling": (C.c_int32, [_VP, C.c_int32]),
    "b200reg_ctx_profile_read": (C.c_int32, [_VP, _VP, C.c_int32, C.POINTER(C.c_int32)]),
    "b200reg_ctx_set_trace": (C.c_int32, [_VP, C.c_int32]),
    "b200reg_ctx_trace_read": (C.c_int32, [_VP, _VP, C.c_int32, C.POINTER(C.c_int32)]),
    "b200reg_cloud_upload_f32": (C.c_int32, [_VP, _VP, C.c_int32, _VP, C.c_int64, C.POINTER(_VP)]),
    "b200reg_cloud_update_f32": (C.c_int32, [_VP, _VP, _VP, C.c_int32, _VP, C.c_int64, C.c_int32]),
    "b200reg_cloud_upload_f64": (C.c_int32, [_VP, _VP, _VP, C.c_int64, C.POINTER(_VP)]),
    "b200reg_cloud_download_f32": (C.c_int32, [_VP, _VP, _VP, C.c_int32, _VP]),
    "b200reg_cloud_size": (C.c_int64, [_VP]),
    "b200reg_cloud_has_normals": (C.c_int32, [_VP]),
    "b200reg_cloud_free": (None, [_VP, _VP]),
    "b200reg_cloud_transform": (C.c_int32, [_VP, _VP, _VP]),
    "b200reg_index_build": (C.c_int32, [_VP, _VP, C.c_float, C.POINTER(_VP)]),
    "b200reg_index_free": (None, [_VP, _VP]),
    "b200reg_index_size": (C.c_int64, [_VP]),
    "b200reg_index_cell": (C.c_float, [_VP]),
    "b200reg_knn": (C.c_int32, [_VP, _VP, _VP, _VP, C.c_int32, C.c_float, C.c_float, C.c_int32, _VP, _VP, C.POINTER(C.c_uint64)]),
    "b200reg_icp_step": (C.c_int32, [_VP, _VP, _VP, _VP, C.c_int32, C.c_float, C.c_float, C.c_float, C.c_int32, _VP, _VP, C.POINTER(StepStats)]),
    "b200reg_trimmed_dist": (C.c_int32, [_VP, _VP, C.c_int64, C.c_float, C.POINTER(C.c_float), _VP]),
    "b200reg_dist_outlier": (C.c_int32, [_VP, C.c_int32, _VP, C.c_int64, C.c_float, C.POINTER(C.c_float), _VP]),
    "b200reg_surface_normal_outlier": (C.c_int32, [_VP, _VP, _VP, C.c_int64, _VP, C.c_int32, C.c_int64, C.c_float, _VP]),
    "b200reg_point_to_plane": (C.c_int32, [_VP, _VP, _VP, _VP, _VP, _VP, C.c_int32, C.c_int32, _VP, _VP, C.POINTER(StepStats)]),
    "b200reg_point_to_plane_residual": (C.c_int32, [_VP, _VP, _VP, _VP, _VP, _VP, C.c_int32, C.c_int32, C.POINTER(C.c_double), C.POINTER(C.c_int64)]),
    "b200reg_point_to_plane_residual_noalloc": (C.c_int32, [_VP, _VP, _VP, _VP, _VP, _VP, C.c_int32, C.c_int32, C.POINTER(C.c_double), C.POINTER(C.c_int64)]),
    "b200reg_solve_point_to_plane": (C.c_int32, [_VP, _VP, _VP, _VP, C.c_int32, C.c_int32, _VP, _VP]),
    "b200reg_icp_create": (C.c_int32, [_VP, C.POINTER(IcpConfig), C.POINTER(_VP)]),
    "b200reg_icp_destroy": (None, [_VP]),
    "b200reg_icp_set_reference": (C.c_int32, [_VP, _VP]),
    "b200reg_icp_register": (C.c_int32, [_VP, _VP, _VP, _VP, C.POINTER(IcpStats)]),
    "b200reg_icp_register_batch": (C.c_int32, [_VP, C.POINTER(IcpConfig), C.c_int32, _VP, _VP, _VP, _VP, _VP, _VP]),
    "b200reg_batch_create": (C.c_int32, [_VP, C.POINTER(IcpConfig), C.c_int32, C.POINTER(_VP)]),
    "b200reg_batch_destroy": (None, [_VP]),
    "b200reg_batch_lanes": (C.c_int32, [_VP]),
    "b200reg_batch_set_reference": (C.c_int32, [_VP, _VP]),
    "b200reg_batch_register": (C.c_int32, [_VP, C.c_int32, _VP, _VP, _VP, _VP, _VP, _VP]),
    "b200reg_batch_register_host": (C.c_int32, [_VP, C.c_int32, _VP, _VP, _VP, _VP, _VP, _VP]),
    "b200reg_multi_create": (C.c_int32, [_VP, C.c_int32, C.POINTER(IcpConfig), C.c_int32, C.POINTER(_VP)]),
    "b200reg_multi_destroy": (None, [_VP]),
    "b200reg_multi_devices": (C.c_int32, [_VP]),
    "b200reg_multi_nccl_version": (C.c_int32, [_VP]),
    "b200reg_multi_last_error": (C.c_char_p, [_VP]),
    "b200reg_multi_launch_count": (C.c_uint64, [_VP]),
    "b200reg_multi_register_host": (C.c_int32, [_VP, C.c_int32, _VP, _VP, _VP, _VP, _VP, _VP, C.POINTER(C.c_float)]),
    "b200reg_icp_index": (_VP, [_VP]),
    "b200reg_icp_reference_mean": (C.c_int32, [_VP, _VP]),
    "b200reg_voxelgrid": (C.c_int32, [_VP, _VP, C.c_float, C.c_float, C.c_float, C.c_int32, C.c_int32]),
    "b200reg_surface_normals": (C.c_int32, [_VP, _VP, C.c_int32, C.c_float, C.c_float, _VP, _VP, _VP, _VP, C.POINTER(C.c_int32)]),
    "b200reg_surface_normals_ex": (C.c_int32, [_VP, _VP, C.POINTER(NormalsOptions), C.POINTER(C.c_int32)]),
    "b200reg_map_create": (C.c_int32, [_VP, C.c_double, C.POINTER(_VP)]),
    "b200reg_map_destroy": (None, [_VP, _VP]),
    "b200reg_map_size": (C.c_int64, [_VP]),
    "b200reg_map_has_normals": (C.c_int32, [_VP]),
    "b200reg_map_insert": (C.c_int32, [_VP, _VP, _VP, _VP, C.c_int64, _VP, C.POINTER(CropSpec)]),
    "b200reg_map_crop": (C.c_int32, [_VP, _VP, C.POINTER(CropSpec), C.POINTER(_VP)]),
    "b200reg_map_download": (C.c_int32, [_VP, _VP, _VP, _VP]),
    "b200reg_voxel_down_sample_f64": (C.c_int32, [_VP, C.c_double, _VP, _VP, C.c_int64, _VP, _VP, C.POINTER(C.c_int64)]),
    "b200reg_crop_f64": (C.c_int32, [_VP, C.POINTER(CropSpec), _VP, _VP, C.c_int64, _VP, _VP, C.POINTER(C.c_int64)]),
    "b200reg_cloud64_upload": (C.c_int32, [_VP, _VP, _VP, C.c_int64, C.POINTER(_VP)]),
    "b200reg_cloud64_prefetch": (C.c_int32, [_VP, _VP, _VP, C.c_int64, C.POINTER(_VP)]),
    "b200reg_cloud64_acquire": (C.c_int32, [_VP, _VP]),
    "b200reg_cloud64_download": (C.c_int32, [_VP, _VP, _VP, _VP]),
    "b200reg_cloud64_size": (C.c_int64, [_VP]),
    "b200reg_cloud64_has_normals": (C.c_int32, [_VP]),
    "b200reg_cloud64_free": (None, [_VP, _VP]),
    "b200reg_cloud64_crop": (C.c_int32, [_VP, C.POINTER(CropSpec), _VP, C.POINTER(_VP)]),
    "b200reg_cloud64_voxel_down_sample": (C.c_int32, [_VP, C.c_double, _VP, C.POINTER(_VP)]),
    "b200reg_cloud64_estimate_normals": (C.c_int32, [_VP, _VP, C.c_int32, C.c_double, _VP, C.POINTER(C.c_int32)]),
    "b200reg_cloud64_to_f32": (C.c_int32, [_VP, _VP, C.POINTER(_VP)]),
    "b200reg_map_insert_cloud64": (C.c_int32, [_VP, _VP, _VP, _VP, C.POINTER(CropSpec)]),
    "b200reg_cloud_from_pointcloud2": (C.c_int32, [_VP, _VP, C.c_size_t, C.POINTER(Pc2Layout), C.POINTER(_VP)]),
    "b200reg_cloud64_from_pointcloud2": (C.c_int32, [_VP, _VP, C.c_size_t, C.POINTER(Pc2Layout), C.POINTER(_VP)]),
    "b200reg_overlap_indices": (C.c_int32, [_VP, _VP, _VP, _VP, C.c_double, C.c_int64, _VP, C.POINTER(C.c_int64), _VP, C.POINTER(C.c_int64)]),
    "b200reg_information_matrix": (C.c_int32, [_VP, _VP, _VP, C.c_double, _VP, _VP, C.POINTER(C.c_int64)]),
    "b200reg_var_trimmed_dist": (C.c_int32, [_VP, _VP, C.c_int64, C.c_float, C.c_float, C.c_float, C.POINTER(C.c_float), C.POINTER(C.c_float), _VP]),
    "b200reg_point_to_point": (C.c_int32, [_VP, _VP, _VP, _VP, _VP, _VP, C.c_int32, _VP, C.POINTER(StepStats)]),
    "b200reg_point_to_plane_covariance": (C.c_int32, [_VP, _VP, _VP, _VP, _VP, _VP, C.c_int32, _VP, _VP, _VP, C.c_float, _VP]),
    "b200reg_cloud_filter": (C.c_int32, [_VP, _VP, C.POINTER(FilterParams)]),
    "b200reg_map_carve": (C.c_int32, [_VP, _VP, _VP, C.c_int64, _VP, C.POINTER(CropSpec), C.POINTER(CarveParams), C.POINTER(C.c_int64)]),
    "b200reg_map_carve_cloud64": (C.c_int32, [_VP, _VP, _VP, _VP, C.POINTER(CropSpec), C.POINTER(CarveParams), C.POINTER(C.c_int64)]),
}


def lib():
    """Load libb200reg.so; fails loudly when the CUDA extension has not been built."""
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                              "(there is no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(L, name)   # AttributeError if the library does not export a declared symbol
            fn.restype = res
            fn.argtypes = args
        _LIB = L
    return _LIB


def _p(a):
    return None if a is None else a.ctypes.data_as(_VP)


def _f32(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float32)


def _f64(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float64)


def crop_spec(kind, center, rmin=0.0, rmax=0.0, zmin=0.0, zmax=0.0, invert=False):
    s = CropSpec(kind=kind, invert=int(invert), rmin=rmin, rmax=rmax, zmin=zmin, zmax=zmax)
    for i in range(3):
        s.center[i] = float(center[i])
    return s


def icp_config(**kw):
    """Defaults = the CPU twins' defaults (MatchersImpl.h:97-100, OutlierFiltersImpl.h:137, TransformationCheckersImpl.h)."""
    c = IcpConfig(knn=1, max_dist=float("inf"), epsilon=0.0, trim_ratio=0.85, max_angle=-1.0, force2D=0, force4DOF=0,
                  max_iter=40, min_diff_rot=0.001, min_diff_trans=0.01, smooth_length=3, bound_rot=-1.0, bound_trans=-1.0,
                  cell=0.0)
    for k, v in kw.items():
        if not hasattr(c, k):
            raise AttributeError(k)
        if k == "checker_order":
            v = (C.c_int32 * 3)(*(list(v) + [0, 0, 0])[:3])
        setattr(c, k, v)
    return c


NOALLOC_MAX_PAIRS = 1 << 20   # B200REG_NOALLOC_MAX_PAIRS


class Context:
    def __init__(self, device=0):
        self.h = _VP()
        code = lib().b200reg_ctx_create(device, C.byref(self.h))
        if code != OK:
            self.h = None
            raise _EXC.get(code, B200RegError)(f"b200reg_ctx_create(device={device}) failed with {code}: no usable CUDA device "
                                               "(this library has no CPU fallback)")

    def check(self, code):
        if code != OK:
            msg = lib().b200reg_last_error(self.h)
            raise _EXC.get(code, B200RegError)((msg or b"").decode("utf-8", "replace"))

    def sync(self):
        self.check(lib().b200reg_ctx_sync(self.h))

    def timer_start(self):
        self.check(lib().b200reg_ctx_timer_start(self.h))

    def timer_stop(self):
        ms = C.c_float(0)
        self.check(lib().b200reg_ctx_timer_stop(self.h, C.byref(ms)))
        return ms.value

    def flush_l2(self):
        self.check(lib().b200reg_ctx_flush_l2(self.h))

    def set_profiling(self, on=True):
        """Bracket every match-kernel launch with its own CUDA event pair (kernel-only time for the roofline)."""
        self.check(lib().b200reg_ctx_set_profiling(self.h, int(on)))

    def profile_read(self, cap=256):
        ms = np.zeros(cap, dtype=np.float32)
        cnt = C.c_int32(0)
        self.check(lib().b200reg_ctx_profile_read(self.h, _p(ms), cap, C.byref(cnt)))
        return ms[:cnt.value].copy()

    def set_trace(self, on=True):
        self.check(lib().b200reg_ctx_set_trace(self.h, int(on)))

    def trace_read(self, cap=2047):
        """[(tag, t_ns)] stamped by the kernels of the last graph-resident registration."""
        buf = np.zeros(cap, dtype=np.uint64)
        cnt = C.c_int32(0)
        self.check(lib().b200reg_ctx_trace_read(self.h, _p(buf), cap, C.byref(cnt)))
        v = buf[:cnt.value]
        return [(int(x >> np.uint64(56)), int(x & np.uint64(0x00ffffffffffffff))) for x in v]

    @property
    def launch_count(self):
        return int(lib().b200reg_ctx_launch_count(self.h))

    def close(self):
        if self.h:
            lib().b200reg_ctx_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- clouds
    def upload(self, feat, normals=None):
        """feat: N x stride float32 (stride >= 3, xyz first) or float64 N x 3 (open3d side)."""
        feat = np.asarray(feat)
        if feat.dtype == np.float64:
            xyz, nrm = _f64(feat[:, :3]), _f64(normals)
            h = _VP()
            self.check(lib().b200reg_cloud_upload_f64(self.h, _p(xyz), _p(nrm), xyz.shape[0], C.byref(h)))
        else:
            feat, nrm = _f32(feat), _f32(normals)
            h = _VP()
            self.check(lib().b200reg_cloud_upload_f32(self.h, _p(feat), feat.shape[1], _p(nrm), feat.shape[0], C.byref(h)))
        return Cloud(self, h)

    def cloud_from_pointcloud2(self, data, layout, out=None):
        """RosPointCloud2Deserializer<float>::deserialize on the device: message bytes -> fp32 cloud (x y z pad, normals)."""
        data = np.ascontiguousarray(np.frombuffer(data, dtype=np.uint8) if not isinstance(data, np.ndarray) else data.view(np.uint8).reshape(-1))
        c = out if out is not None else Cloud(self, _VP())
        self.check(lib().b200reg_cloud_from_pointcloud2(self.h, _p(data) if data.size else None, data.size, C.byref(layout), C.byref(c.h)))
        return c

    def cloud64_from_pointcloud2(self, data, layout, out=None):
        """open3d_conversions::rosToOpen3d on the device: message bytes -> fp64 cloud (points, normals)."""
        data = np.ascontiguousarray(np.frombuffer(data, dtype=np.uint8) if not isinstance(data, np.ndarray) else data.view(np.uint8).reshape(-1))
        c = out if out is not None else Cloud64(self)
        self.check(lib().b200reg_cloud64_from_pointcloud2(self.h, _p(data) if data.size else None, data.size, C.byref(layout), C.byref(c.h)))
        return c

    def overlap_indices(self, source, target, T_source_to_target, voxel_size, min_points_per_voxel=1):
        """computeIndicesOfOverlappingPoints on two Cloud64: ascending indices (source, target) of the points in shared voxels."""
        T = _f64(T_source_to_target).reshape(16)
        si, ti = np.empty(source.size, dtype=np.int64), np.empty(target.size, dtype=np.int64)
        ns, nt = C.c_int64(0), C.c_int64(0)
        self.check(lib().b200reg_overlap_indices(self.h, source.h, target.h, _p(T), float(voxel_size), int(min_points_per_voxel),
                                                 _p(si), C.byref(ns), _p(ti), C.byref(nt)))
        return si[:ns.value], ti[:nt.value]

    def information_matrix(self, source, target, max_correspondence_distance, T_source_to_target):
        """GetInformationMatrixFromPointClouds on two Cloud64: (6 x 6 fp64, number of correspondences)."""
        T = _f64(T_source_to_target).reshape(16)
        info = np.zeros(36, dtype=np.float64)
        nc = C.c_int64(0)
        self.check(lib().b200reg_information_matrix(self.h, source.h, target.h, float(max_correspondence_distance), _p(T), _p(info), C.byref(nc)))
        return info.reshape(6, 6), nc.value

    def build_index(self, cloud, cell=0.0):
        h = _VP()
        self.check(lib().b200reg_index_build(self.h, cloud.h, cell, C.byref(h)))
        return Index(self, h)

    def knn(self, index, query, T=None, k=1, max_dist=np.inf, epsilon=0.0, allow_self=True, want=True):
        n = query.size
        ids = np.empty((n, k), dtype=np.int32) if want else None
        d2 = np.empty((n, k), dtype=np.float32) if want else None
        visits = C.c_uint64(0)
        Tm = None if T is None else _f32(T).reshape(16)
        self.check(lib().b200reg_knn(self.h, index.h, query.h, _p(Tm), k, max_dist, epsilon, int(allow_self), _p(ids), _p(d2), C.byref(visits)))
        return ids, d2, visits.value

    def icp_step(self, index, reading, T=None, k=1, max_dist=np.inf, trim_ratio=-1.0, max_angle=-1.0, force2D=False):
        A = np.zeros((6, 6), dtype=np.float64)
        b = np.zeros(6, dtype=np.float64)
        st = StepStats()
        Tm = None if T is None else _f32(T).reshape(16)
        self.check(lib().b200reg_icp_step(self.h, index.h, reading.h, _p(Tm), k, max_dist, trim_ratio, max_angle, int(force2D), _p(A), _p(b), C.byref(st)))
        return A, b, st

    def trimmed_dist(self, d2, ratio):
        d2 = _f32(d2)
        tau = C.c_float(0)
        w = np.empty(d2.shape, dtype=np.float32)
        self.check(lib().b200reg_trimmed_dist(self.h, _p(d2), d2.size, ratio, C.byref(tau), _p(w)))
        return tau.value, w

    def surface_normal_outlier(self, nrm_read, nrm_ref, ids, max_angle):
        ids = np.ascontiguousarray(ids, dtype=np.int32)
        n, k = ids.shape
        nr, nf = _f32(nrm_read), _f32(nrm_ref)
        w = np.empty((n, k), dtype=np.float32)
        self.check(lib().b200reg_surface_normal_outlier(self.h, _p(nr), _p(nf), 0 if nf is None else nf.shape[0], _p(ids), k, n, max_angle, _p(w)))
        return w

    def point_to_plane(self, reading, reference, ids, d2, w, force2D=False):
        ids = np.ascontiguousarray(ids, dtype=np.int32)
        d2, w = _f32(d2), _f32(w)
        k = ids.size // max(reading.size, 1)
        A = np.zeros((6, 6), dtype=np.float64)
        b = np.zeros(6, dtype=np.float64)
        st = StepStats()
        self.check(lib().b200reg_point_to_plane(self.h, reading.h, reference.h, _p(ids), _p(d2), _p(w), k, int(force2D), _p(A), _p(b), C.byref(st)))
        return A, b, st

    def point_to_plane_residual(self, reading, reference, ids, d2, w, force2D=False):
        """PointToPlaneErrorMinimizer::getResidualError on explicit matches / weights: (sum w ((p - q) . n)^2, kept pairs)."""
        ids = np.ascontiguousarray(ids, dtype=np.int32)
        d2, w = _f32(d2), _f32(w)
        k = ids.size // max(reading.size, 1)
        res, kept = C.c_double(0), C.c_int64(0)
        self.check(lib().b200reg_point_to_plane_residual(self.h, reading.h, reference.h, _p(ids), _p(d2), _p(w), k, int(force2D), C.byref(res), C.byref(kept)))
        return res.value, kept.value

    def point_to_plane_residual_noalloc(self, reading, reference, ids, d2, w, force2D=False):
        """The allocation-free variant (static device scratch, at most NOALLOC_MAX_PAIRS pairs); same result as point_to_plane_residual."""
        ids = np.ascontiguousarray(ids, dtype=np.int32)
        d2, w = _f32(d2), _f32(w)
        k = ids.size // max(reading.size, 1)
        res, kept = C.c_double(0), C.c_int64(0)
        self.check(lib().b200reg_point_to_plane_residual_noalloc(self.h, reading.h, reference.h, _p(ids), _p(d2), _p(w), k, int(force2D),
                                                                 C.byref(res), C.byref(kept)))
        return res.value, kept.value

    def voxel_down_sample(self, voxel, xyz, nrm=None):
        """o3d_slam::voxelize / Open3D VoxelDownSample on a host fp64 cloud (ascending key order)."""
        xyz, nrm = _f64(xyz), _f64(nrm)
        xo = np.empty_like(xyz)
        no = None if nrm is None else np.empty_like(nrm)
        m = C.c_int64(0)
        self.check(lib().b200reg_voxel_down_sample_f64(self.h, float(voxel), _p(xyz), _p(nrm), xyz.shape[0], _p(xo), _p(no), C.byref(m)))
        return xo[:m.value], (None if no is None else no[:m.value])

    def dist_outlier(self, kind, d2, param):
        """MaxDist (0) / MinDist (1) / MedianDist (2) outlier weights on squared distances; returns (limit, weights)."""
        d2 = _f32(d2)
        w = np.empty(d2.shape, dtype=np.float32)
        lim = C.c_float(0)
        self.check(lib().b200reg_dist_outlier(self.h, int(kind), _p(d2), d2.size, float(param), C.byref(lim), _p(w)))
        return lim.value, w

    def var_trimmed_dist(self, d2, min_ratio, max_ratio, lam):
        """VarTrimmedDistOutlierFilter on squared distances; returns (optimised ratio, limit, weights)."""
        d2 = _f32(d2)
        w = np.empty(d2.shape, dtype=np.float32)
        ratio, tau = C.c_float(0), C.c_float(0)
        self.check(lib().b200reg_var_trimmed_dist(self.h, _p(d2), d2.size, float(min_ratio), float(max_ratio), float(lam), C.byref(ratio),
                                                  C.byref(tau), _p(w)))
        return ratio.value, tau.value, w

    def point_to_point(self, reading, reference, ids, d2, w):
        """PointToPointErrorMinimizer::compute on explicit matches / weights; returns (T 4x4 float32, StepStats)."""
        ids = np.ascontiguousarray(ids, dtype=np.int32)
        d2, w = _f32(d2), _f32(w)
        k = ids.size // max(reading.size, 1)
        T = np.zeros((4, 4), dtype=np.float32)
        st = StepStats()
        self.check(lib().b200reg_point_to_point(self.h, reading.h, reference.h, _p(ids), _p(d2), _p(w), k, _p(T), C.byref(st)))
        return T, st

    def point_to_plane_covariance(self, reading, reference, ids, d2, w, mean_reading, mean_reference, T, sensor_std_dev):
        """PointToPlaneWithCovErrorMinimizer::estimateCovariance for the step transform T; returns the 6x6 covariance."""
        ids = np.ascontiguousarray(ids, dtype=np.int32)
        d2, w, T = _f32(d2), _f32(w), _f32(T)
        mr, ms = _f64(mean_reading), _f64(mean_reference)
        k = ids.size // max(reading.size, 1)
        cov = np.zeros((6, 6), dtype=np.float64)
        self.check(lib().b200reg_point_to_plane_covariance(self.h, reading.h, reference.h, _p(ids), _p(d2), _p(w), k, _p(mr), _p(ms), _p(T),
                                                           float(sensor_std_dev), _p(cov)))
        return cov

    def crop_f64(self, spec, xyz, nrm=None):
        xyz, nrm = _f64(xyz), _f64(nrm)
        xo = np.empty_like(xyz)
        no = None if nrm is None else np.empty_like(nrm)
        m = C.c_int64(0)
        self.check(lib().b200reg_crop_f64(self.h, C.byref(spec), _p(xyz), _p(nrm), xyz.shape[0], _p(xo), _p(no), C.byref(m)))
        return xo[:m.value], (None if no is None else no[:m.value])


def solve_point_to_plane(A, b, mean_reading, mean_reference, force2D=False, force4DOF=False):
    A, b = _f64(A).reshape(36), _f64(b)
    mr, ms = _f64(mean_reading), _f64(mean_reference)
    T = np.empty(16, dtype=np.float32)
    x = np.zeros(6, dtype=np.float64)
    code = lib().b200reg_solve_point_to_plane(_p(A), _p(b), _p(mr), _p(ms), int(force2D), int(force4DOF), _p(T), _p(x))
    if code != OK:
        raise _EXC.get(code, B200RegError)("solve_point_to_plane failed")
    return T.reshape(4, 4), x


class Cloud:
    def __init__(self, ctx, h):
        self.ctx, self.h = ctx, h

    @property
    def size(self):
        return int(lib().b200reg_cloud_size(self.h))

    @property
    def has_normals(self):
        return bool(lib().b200reg_cloud_has_normals(self.h))

    def update(self, feat, normals=None, wait=True):
        """Re-fill this device cloud from host arrays.  wait=False only enqueues the copy: keep `feat` alive and untouched
        until the next result-returning call (use pinned memory, e.g. torch's pin_memory, for a truly asynchronous copy)."""
        feat, nrm = _f32(feat), _f32(normals)
        self._keep = (feat, nrm)
        self.ctx.check(lib().b200reg_cloud_update_f32(self.ctx.h, self.h, _p(feat), feat.shape[1], _p(nrm), feat.shape[0], 0 if wait else 1))

    def download(self, stride=4, normals=None):
        n = self.size
        feat = np.empty((n, stride), dtype=np.float32)
        want_n = self.has_normals if normals is None else normals
        nrm = np.empty((n, 3), dtype=np.float32) if want_n else None
        self.ctx.check(lib().b200reg_cloud_download_f32(self.ctx.h, self.h, _p(feat), stride, _p(nrm)))
        return feat, nrm

    def transform(self, T):
        T = _f32(T).reshape(16)
        self.ctx.check(lib().b200reg_cloud_transform(self.ctx.h, self.h, _p(T)))

    def filter(self, kind, dim=-1, dist=1.0, box=(-1, 1, -1, 1, -1, 1), remove_inside=True, random_method=0, prob=0.75, seed=1):
        """MaxDist / MinDist / BoundingBox / RandomSampling DataPointsFilter, in place (order preserving)."""
        P = FilterParams(kind=int(kind), dim=int(dim), dist=float(dist), box=(C.c_float * 6)(*[float(b) for b in box]),
                         remove_inside=int(remove_inside), random_method=int(random_method), prob=float(prob), seed=int(seed))
        self.ctx.check(lib().b200reg_cloud_filter(self.ctx.h, self.h, C.byref(P)))
        return self

    def voxelgrid(self, vx, vy, vz, use_centroid=True, average_descriptors=True):
        self.ctx.check(lib().b200reg_voxelgrid(self.ctx.h, self.h, vx, vy, vz, int(use_centroid), int(average_descriptors)))

    def surface_normals(self, knn, max_dist=np.inf, epsilon=0.0, orient_origin=None, keep=(), smooth=False):
        """SurfaceNormalDataPointsFilter on this cloud (writes its normals).  keep: any of densities, eig_values, eig_vectors, matched_ids,
        mean_dists, lin_plan_curv."""
        n = self.size
        oo = None if orient_origin is None else _f32(orient_origin)
        shapes = dict(densities=(n,), eig_values=(n, 3), eig_vectors=(n, 9), matched_ids=(n, knn), mean_dists=(n,), lin_plan_curv=(n, 3))
        out = {k: (np.empty(sh, dtype=np.float32) if k in keep else None) for k, sh in shapes.items()}
        opt = NormalsOptions(knn=int(knn), max_dist=float(max_dist), epsilon=float(epsilon), smooth_normals=int(smooth),
                             orient_origin=None if oo is None else oo.ctypes.data, **{k: (None if v is None else v.ctypes.data) for k, v in out.items()})
        deg = C.c_int32(0)
        self.ctx.check(lib().b200reg_surface_normals_ex(self.ctx.h, self.h, C.byref(opt), C.byref(deg)))
        out["degenerate"] = deg.value
        return out

    def free(self):
        if self.h:
            lib().b200reg_cloud_free(self.ctx.h, self.h)
            self.h = None

    def __del__(self):
        try:
            if self.ctx.h:
                self.free()
        except Exception:
            pass


class Index:
    def __init__(self, ctx, h, owned=True):
        self.ctx, self.h, self.owned = ctx, h, owned

    @property
    def size(self):
        return int(lib().b200reg_index_size(self.h))

    @property
    def cell(self):
        return float(lib().b200reg_index_cell(self.h))

    def free(self):
        if self.h and self.owned:
            lib().b200reg_index_free(self.ctx.h, self.h)
        self.h = None

    def __del__(self):
        try:
            if self.ctx.h:
                self.free()
        except Exception:
            pass


class ICP:
    """PointMatcher<float>::ICP with the CUDA chain: KDTree matcher + TrimmedDist/SurfaceNormal filters + PointToPlane."""

    def __init__(self, ctx, cfg=None, **kw):
        self.ctx = ctx
        self.cfg = cfg if cfg is not None else icp_config(**kw)
        self.h = _VP()
        ctx.check(lib().b200reg_icp_create(ctx.h, C.byref(self.cfg), C.byref(self.h)))

    def set_reference(self, cloud):
        self.ctx.check(lib().b200reg_icp_set_reference(self.h, cloud.h))

    @property
    def index(self):
        return Index(self.ctx, lib().b200reg_icp_index(self.h), owned=False)

    @property
    def reference_mean(self):
        m = np.zeros(3, dtype=np.float32)
        lib().b200reg_icp_reference_mean(self.h, _p(m))
        return m

    def register(self, reading, T_init=None):
        T_init = np.eye(4, dtype=np.float32) if T_init is None else _f32(T_init)
        T = np.zeros(16, dtype=np.float32)
        st = IcpStats()
        code = lib().b200reg_icp_register(self.h, reading.h, _p(T_init.reshape(16)), _p(T), C.byref(st))
        self.last_stats = st
        self.ctx.check(code)
        return T.reshape(4, 4), st

    def close(self):
        if self.h:
            lib().b200reg_icp_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            if self.ctx.h:
                self.close()
        except Exception:
            pass


def register_batch(ctx, cfg, refs, readings, T_init):
    """n independent registrations on ctx's device (b200reg_icp_register_batch).  refs / readings: lists of Cloud;
    T_init: n x 4 x 4.  Returns (T n x 4 x 4 float32, [IcpStats], status int32[n])."""
    n = len(refs)
    assert len(readings) == n
    Ti = np.ascontiguousarray(np.asarray(T_init, dtype=np.float32).reshape(n, 16))
    To = np.zeros((n, 16), dtype=np.float32)
    stats = (IcpStats * max(n, 1))()
    status = np.zeros(max(n, 1), dtype=np.int32)
    rh = (_VP * max(n, 1))(*[c.h.value if isinstance(c.h, _VP) else c.h for c in refs])
    qh = (_VP * max(n, 1))(*[c.h.value if isinstance(c.h, _VP) else c.h for c in readings])
    ctx.check(lib().b200reg_icp_register_batch(ctx.h, C.byref(cfg), n, C.cast(rh, _VP), C.cast(qh, _VP), _p(Ti), _p(To), C.cast(stats, _VP),
                                               _p(status)))
    return To.reshape(n, 4, 4), list(stats)[:n], status[:n]


class Batch:
    """b200reg_batch: `lanes` registrations in flight on the ctx's device (pair i -> lane i mod lanes)."""

    def __init__(self, ctx, cfg=None, lanes=0, **kw):
        self.ctx = ctx
        self.cfg = cfg if cfg is not None else icp_config(**kw)
        self.h = _VP()
        ctx.check(lib().b200reg_batch_create(ctx.h, C.byref(self.cfg), int(lanes), C.byref(self.h)))

    @property
    def lanes(self):
        return int(lib().b200reg_batch_lanes(self.h))

    def _out(self, n, T_init):
        Ti = np.ascontiguousarray(np.asarray(T_init, dtype=np.float32).reshape(n, 16))
        return Ti, np.zeros((n, 16), dtype=np.float32), (IcpStats * max(n, 1))(), np.zeros(max(n, 1), dtype=np.int32)

    def set_reference(self, cloud):
        """One reference for every pair registered with refs=None."""
        self.ctx.check(lib().b200reg_batch_set_reference(self.h, cloud.h))

    def register(self, refs, readings, T_init):
        """refs / readings: lists of device Clouds of this ctx (refs=None: the shared reference).
        Returns (T n x 4 x 4, [IcpStats], status int32[n])."""
        n = len(readings)
        assert refs is None or len(refs) == n
        Ti, To, stats, status = self._out(n, T_init)
        rh = None if refs is None else C.cast((_VP * max(n, 1))(*[c.h.value if isinstance(c.h, _VP) else c.h for c in refs]), _VP)
        qh = (_VP * max(n, 1))(*[c.h.value if isinstance(c.h, _VP) else c.h for c in readings])
        self.ctx.check(lib().b200reg_batch_register(self.h, n, rh, C.cast(qh, _VP), _p(Ti), _p(To), C.cast(stats, _VP), _p(status)))
        return To.reshape(n, 4, 4), list(stats)[:n], status[:n]

    @staticmethod
    def host_clouds(items):
        """items: [(feat N x stride float32 C-contiguous, normals N x 3 float32 or None)] -> (HostCloud array, keep-alive list)."""
        arr = (HostCloud * max(len(items), 1))()
        keep = []
        for i, (feat, nrm) in enumerate(items):
            assert feat.dtype == np.float32 and feat.flags.c_contiguous and (nrm is None or (nrm.dtype == np.float32 and nrm.flags.c_contiguous))
            keep.append((feat, nrm))
            arr[i] = HostCloud(feat.ctypes.data, feat.shape[1], None if nrm is None else nrm.ctypes.data, feat.shape[0])
        return arr, keep

    def register_host(self, refs, readings, T_init):
        """refs / readings: HostCloud arrays from host_clouds() (keep the numpy arrays alive).  The uploads are part of the pipeline."""
        n = len(readings[1]) if isinstance(readings, tuple) else len(readings)
        ra = refs[0] if isinstance(refs, tuple) else refs
        qa = readings[0] if isinstance(readings, tuple) else readings
        Ti, To, stats, status = self._out(n, T_init)
        self.ctx.check(lib().b200reg_batch_register_host(self.h, n, None if ra is None else C.cast(ra, _VP), C.cast(qa, _VP), _p(Ti), _p(To),
                                                         C.cast(stats, _VP), _p(status)))
        return To.reshape(n, 4, 4), list(stats)[:n], status[:n]

    def close(self):
        if self.h:
            lib().b200reg_batch_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            if self.ctx.h:
                self.close()
        except Exception:
            pass


class Multi:
    """b200reg_multi: one process driving every GPU of the box (one host thread + one b200reg_batch per device), pair i -> device
    i mod G, one ncclAllGather of the results."""

    def __init__(self, devices=None, cfg=None, lanes=0, **kw):
        self.cfg = cfg if cfg is not None else icp_config(**kw)
        self.h = _VP()
        dv = None if not devices else (C.c_int32 * len(devices))(*devices)
        code = lib().b200reg_multi_create(None if dv is None else C.cast(dv, _VP), 0 if dv is None else len(devices), C.byref(self.cfg), int(lanes),
                                          C.byref(self.h))
        if code != OK:
            msg = lib().b200reg_multi_last_error(self.h).decode("utf-8", "replace") if self.h else "b200reg_multi_create failed (no CUDA device)"
            self.close()
            raise _EXC.get(code, B200RegError)(msg)

    @property
    def devices(self):
        return int(lib().b200reg_multi_devices(self.h))

    @property
    def nccl_version(self):
        return int(lib().b200reg_multi_nccl_version(self.h))

    @property
    def launch_count(self):
        return int(lib().b200reg_multi_launch_count(self.h))

    def register_host(self, refs, readings, T_init):
        """refs / readings: results of Batch.host_clouds().  Returns (T n x 4 x 4, [IcpStats], status, gather_ms)."""
        n = len(readings[1])
        Ti = np.ascontiguousarray(np.asarray(T_init, dtype=np.float32).reshape(n, 16))
        To = np.zeros((n, 16), dtype=np.float32)
        stats = (IcpStats * max(n, 1))()
        status = np.zeros(max(n, 1), dtype=np.int32)
        gms = C.c_float(0)
        code = lib().b200reg_multi_register_host(self.h, n, C.cast(refs[0], _VP), C.cast(readings[0], _VP), _p(Ti), _p(To), C.cast(stats, _VP), _p(status),
                                                 C.byref(gms))
        if code != OK:
            raise _EXC.get(code, B200RegError)(lib().b200reg_multi_last_error(self.h).decode("utf-8", "replace"))
        return To.reshape(n, 4, 4), list(stats)[:n], status[:n], gms.value

    def close(self):
        if self.h:
            lib().b200reg_multi_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Cloud64:
    """Device-resident fp64 cloud (the open3d::geometry::PointCloud mirror): the scan-processing chain of
    ScanToMapRegistration.cpp:36-69 without host round trips.  Every step returns a new Cloud64, or refills `out` when given."""

    def __init__(self, ctx, h=None):
        self.ctx = ctx
        self.h = h if h is not None else _VP()

    @classmethod
    def upload(cls, ctx, xyz, nrm=None, out=None):
        xyz, nrm = _f64(xyz), _f64(nrm)
        c = out if out is not None else cls(ctx)
        ctx.check(lib().b200reg_cloud64_upload(ctx.h, _p(xyz), _p(nrm), xyz.shape[0], C.byref(c.h)))
        return c

    @classmethod
    def prefetch(cls, ctx, xyz, nrm=None, out=None):
        """The upload on the copy stream (b200reg_cloud64_prefetch); xyz / nrm must be C-contiguous float64 arrays (they are kept
        referenced until acquire()).  Call acquire() before any other use."""
        for a in (xyz, nrm):
            if a is not None and not (isinstance(a, np.ndarray) and a.dtype == np.float64 and a.flags.c_contiguous):
                raise ValueError("prefetch needs C-contiguous float64 arrays")
        c = out if out is not None else cls(ctx)
        ctx.check(lib().b200reg_cloud64_prefetch(ctx.h, _p(xyz), _p(nrm), xyz.shape[0], C.byref(c.h)))
        c._pending = (xyz, nrm)
        return c

    def acquire(self):
        self.ctx.check(lib().b200reg_cloud64_acquire(self.ctx.h, self.h))
        self._pending = None
        return self

    @property
    def size(self):
        return int(lib().b200reg_cloud64_size(self.h)) if self.h else 0

    @property
    def has_normals(self):
        return bool(lib().b200reg_cloud64_has_normals(self.h)) if self.h else False

    def download(self):
        n = self.size
        xyz = np.empty((n, 3), dtype=np.float64)
        nrm = np.empty((n, 3), dtype=np.float64) if self.has_normals else None
        self.ctx.check(lib().b200reg_cloud64_download(self.ctx.h, self.h, _p(xyz), _p(nrm)))
        return xyz, nrm

    def crop(self, volume, out=None):
        c = out if out is not None else Cloud64(self.ctx)
        self.ctx.check(lib().b200reg_cloud64_crop(self.ctx.h, C.byref(volume), self.h, C.byref(c.h)))
        return c

    def voxel_down_sample(self, voxel, out=None):
        c = out if out is not None else Cloud64(self.ctx)
        self.ctx.check(lib().b200reg_cloud64_voxel_down_sample(self.ctx.h, float(voxel), self.h, C.byref(c.h)))
        return c

    def estimate_normals(self, knn, radius, origin=None):
        o = None if origin is None else _f64(origin)
        deg = C.c_int32(0)
        self.ctx.check(lib().b200reg_cloud64_estimate_normals(self.ctx.h, self.h, int(knn), float(radius), _p(o), C.byref(deg)))
        return deg.value

    def to_f32(self, out=None):
        """open3dToPointmatcher: the fp32 DataPoints mirror (a capi.Cloud)."""
        c = out if out is not None else Cloud(self.ctx, _VP())
        self.ctx.check(lib().b200reg_cloud64_to_f32(self.ctx.h, self.h, C.byref(c.h)))
        return c

    def free(self):
        if self.h:
            lib().b200reg_cloud64_free(self.ctx.h, self.h)
            self.h = _VP()

    def __del__(self):
        try:
            if self.ctx.h:
                self.free()
        except Exception:
            pass


class VoxelMap:
    """Resident fp64 voxel map (open3d_slam Submap::mapCloud_ + insertScan + cropSubmap)."""

    def __init__(self, ctx, voxel_size):
        self.ctx = ctx
        self.h = _VP()
        ctx.check(lib().b200reg_map_create(ctx.h, voxel_size, C.byref(self.h)))

    @property
    def size(self):
        return int(lib().b200reg_map_size(self.h))

    def insert(self, scan_xyz, scan_nrm, T_map_scan, volume):
        xyz, nrm = _f64(scan_xyz), _f64(scan_nrm)
        T = _f64(T_map_scan).reshape(16)
        self.ctx.check(lib().b200reg_map_insert(self.ctx.h, self.h, _p(xyz), _p(nrm), xyz.shape[0], _p(T), C.byref(volume)))

    def carve(self, scan, T_map_scan, volume, voxel_size=0.1, max_raytracing_length=20.0, truncation_distance=0.1, min_dot=0.5):
        """Space carving (Submap::carve): scan = raw scan in the sensor frame, host fp64 array or a Cloud64.  Returns the removed count."""
        T = _f64(T_map_scan).reshape(16)
        cp = CarveParams(voxel_size, max_raytracing_length, truncation_distance, min_dot)
        removed = C.c_int64(0)
        if isinstance(scan, Cloud64):
            self.ctx.check(lib().b200reg_map_carve_cloud64(self.ctx.h, self.h, scan.h, _p(T), C.byref(volume), C.byref(cp), C.byref(removed)))
        else:
            xyz = _f64(scan)
            self.ctx.check(lib().b200reg_map_carve(self.ctx.h, self.h, _p(xyz), xyz.shape[0], _p(T), C.byref(volume), C.byref(cp), C.byref(removed)))
        return removed.value

    def insert_cloud64(self, scan, T_map_scan, volume):
        T = _f64(T_map_scan).reshape(16)
        self.ctx.check(lib().b200reg_map_insert_cloud64(self.ctx.h, self.h, scan.h, _p(T), C.byref(volume)))

    def crop(self, volume):
        h = _VP()
        self.ctx.check(lib().b200reg_map_crop(self.ctx.h, self.h, C.byref(volume), C.byref(h)))
        return Cloud(self.ctx, h)

    def download(self):
        n = self.size
        xyz = np.empty((n, 3), dtype=np.float64)
        nrm = np.empty((n, 3), dtype=np.float64) if lib().b200reg_map_has_normals(self.h) else None
        self.ctx.check(lib().b200reg_map_download(self.ctx.h, self.h, _p(xyz), _p(nrm)))
        return xyz, nrm

    def close(self):
        if self.h:
            lib().b200reg_map_destroy(self.ctx.h, self.h)
            self.h = None

    def __del__(self):
        try:
            if self.ctx.h:
                self.close()
        except Exception:
            pass
