"""ctypes window onto the C++ scan-to-map Mapper driver (host/mapper.h — the control flow of o3d_slam::Mapper::addRangeMeasurement,
open3d_slam/src/Mapper.cpp:168-504, over the C ABI).  Used by the mapping-loop tests and the config #3 harness."""
import ctypes as C

import numpy as np

from . import capi, pm_shim


class MapperParams(C.Structure):
    _fields_ = [("map_builder_cropper", capi.CropSpec), ("scan_matcher_cropper", capi.CropSpec), ("scan_voxel_size", C.c_double),
                ("normals_radius", C.c_double), ("map_voxel_size", C.c_double), ("min_movement", C.c_double),
                ("reference_period", C.c_double), ("normals_knn", C.c_int32), ("pad", C.c_int32), ("icp", capi.IcpConfig),
                ("carving_enabled", C.c_int32), ("carve_every", C.c_int32), ("carving", capi.CarveParams)]


class MapperStats(C.Structure):
    _fields_ = [("iterations", C.c_int32), ("icp_status", C.c_int32), ("reference_refreshed", C.c_int32), ("inserted", C.c_int32),
                ("reading_points", C.c_int64), ("merge_points", C.c_int64), ("patch_points", C.c_int64), ("map_points", C.c_int64),
                ("stage_ms", C.c_double * 9), ("carved_points", C.c_int64),
                ("prefetch_wait_ms", C.c_double), ("prefetched", C.c_int64)]

    STAGES = ("wide_crop", "voxelize", "normals", "narrow_crop", "upload", "patch_crop", "init_reference", "register", "insert")


def _lib():
    L = pm_shim.lib()
    if not getattr(L, "_b200map_ready", False):
        L.b200map_create.restype = C.c_void_p
        L.b200map_create.argtypes = [C.c_void_p, C.POINTER(MapperParams)]
        L.b200map_destroy.argtypes = [C.c_void_p]
        L.b200map_set_pose.argtypes = [C.c_void_p, C.c_void_p]
        L.b200map_add_range_measurement.restype = C.c_int32
        L.b200map_add_range_measurement.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_double, C.c_void_p, C.c_void_p,
                                                    C.POINTER(MapperStats), C.c_char_p, C.c_int32]
        L.b200map_add_pointcloud2.restype = C.c_int32
        L.b200map_add_pointcloud2.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.POINTER(capi.Pc2Layout), C.c_double, C.c_void_p, C.c_void_p,
                                              C.POINTER(MapperStats), C.c_char_p, C.c_int32]
        L.b200map_prefetch.restype = C.c_int32
        L.b200map_prefetch.argtypes = [C.c_void_p, C.c_void_p, C.c_int64]
        L.b200map_map.restype = C.c_void_p
        L.b200map_map.argtypes = [C.c_void_p]
        L._b200map_ready = True
    return L


class Mapper:
    """One active submap, no carving: see host/mapper.h for what is mirrored and what is left out."""

    def __init__(self, ctx, map_builder_cropper, scan_matcher_cropper, icp_cfg, scan_voxel_size=0.1, normals_knn=20, normals_radius=1.0,
                 map_voxel_size=0.15, min_movement=0.0, reference_period=2.0, carving=None, carve_every=10):
        """carving: None (off, the reference's default) or a capi.CarveParams."""
        self.ctx = ctx
        self.params = MapperParams(map_builder_cropper=map_builder_cropper, scan_matcher_cropper=scan_matcher_cropper,
                                   scan_voxel_size=scan_voxel_size, normals_radius=normals_radius, map_voxel_size=map_voxel_size,
                                   min_movement=min_movement, reference_period=reference_period, normals_knn=normals_knn, pad=0, icp=icp_cfg,
                                   carving_enabled=0 if carving is None else 1, carve_every=carve_every,
                                   carving=carving if carving is not None else capi.CarveParams(0.1, 20.0, 0.1, 0.5))
        self.h = _lib().b200map_create(ctx.h, C.byref(self.params))
        if not self.h:
            raise RuntimeError("b200map_create failed")

    def set_pose(self, T):
        T = np.ascontiguousarray(T, dtype=np.float64).reshape(16)
        _lib().b200map_set_pose(self.h, T.ctypes.data_as(C.c_void_p))

    def add_range_measurement(self, xyz, time, odom, normals=None):
        xyz = np.ascontiguousarray(xyz, dtype=np.float64)
        nrm = None if normals is None else np.ascontiguousarray(normals, dtype=np.float64)
        odom = np.ascontiguousarray(odom, dtype=np.float64).reshape(16)
        pose = np.zeros(16, dtype=np.float64)
        st = MapperStats()
        err = C.create_string_buffer(1024)
        ok = _lib().b200map_add_range_measurement(self.h, xyz.ctypes.data_as(C.c_void_p), None if nrm is None else nrm.ctypes.data_as(C.c_void_p),
                                                  xyz.shape[0], float(time), odom.ctypes.data_as(C.c_void_p), pose.ctypes.data_as(C.c_void_p),
                                                  C.byref(st), err, len(err))
        return bool(ok), pose.reshape(4, 4), st, err.value.decode("utf-8", "replace")

    def prefetch(self, xyz):
        """Hand a COMING scan (a C-contiguous float64 n x 3 array) to the Mapper's worker, which uploads it and runs the scan-processing
        chain ahead while this thread registers the current scan; the array must stay alive and unchanged until the
        add_range_measurement(xyz, ...) call that passes the SAME array."""
        if not (isinstance(xyz, np.ndarray) and xyz.dtype == np.float64 and xyz.flags.c_contiguous):
            raise ValueError("prefetch needs a C-contiguous float64 array (the same object is passed to add_range_measurement)")
        if not _lib().b200map_prefetch(self.h, xyz.ctypes.data_as(C.c_void_p), xyz.shape[0]):
            raise RuntimeError("b200map_prefetch failed")

    def add_pointcloud2(self, data, layout, time, odom):
        """The same step fed with the bytes of a sensor_msgs/PointCloud2 (layout: capi.pc2_layout)."""
        data = np.ascontiguousarray(np.frombuffer(data, dtype=np.uint8) if not isinstance(data, np.ndarray) else data.view(np.uint8).reshape(-1))
        odom = np.ascontiguousarray(odom, dtype=np.float64).reshape(16)
        pose = np.zeros(16, dtype=np.float64)
        st = MapperStats()
        err = C.create_string_buffer(1024)
        ok = _lib().b200map_add_pointcloud2(self.h, data.ctypes.data_as(C.c_void_p), data.size, C.byref(layout), float(time),
                                            odom.ctypes.data_as(C.c_void_p), pose.ctypes.data_as(C.c_void_p), C.byref(st), err, len(err))
        return bool(ok), pose.reshape(4, 4), st, err.value.decode("utf-8", "replace")

    def map_points(self):
        """Download the resident map (fp64 points and normals)."""
        mh = _lib().b200map_map(self.h)
        n = int(capi.lib().b200reg_map_size(mh))
        xyz = np.zeros((n, 3), dtype=np.float64)
        nrm = np.zeros((n, 3), dtype=np.float64) if capi.lib().b200reg_map_has_normals(mh) else None
        self.ctx.check(capi.lib().b200reg_map_download(self.ctx.h, mh, xyz.ctypes.data_as(C.c_void_p),
                                                       None if nrm is None else nrm.ctypes.data_as(C.c_void_p)))
        return xyz, nrm

    def close(self):
        if self.h:
            _lib().b200map_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            if self.ctx.h:
                self.close()
        except Exception:
            pass
