"""ctypes window onto libb200pm.so — the C++ shim that mirrors libpointmatcher's registrable-module surface
(host/pm_compat.h, host/cuda_modules.h).  Lets Python drive a PM::ICP built from the reference's YAML chain files with the
Cuda* modules, and catches the libpointmatcher exception types by name."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libb200pm.so")
_LIB = None

KINDS = {"DataPointsFilter": 0, "Matcher": 1, "OutlierFilter": 2, "ErrorMinimizer": 3, "TransformationChecker": 4,
         "Transformation": 5, "Inspector": 6, "Logger": 7}


class PmError(RuntimeError):
    pass


class InvalidParameter(PmError):
    pass


class InvalidElement(PmError):
    pass


class InvalidModuleType(PmError):
    pass


class ConvergenceError(PmError):
    pass


class TransformationError(PmError):
    pass


class InvalidField(PmError):
    pass


class ConfigurationError(PmError):
    pass


_EXC = {1: InvalidParameter, 2: InvalidElement, 3: InvalidModuleType, 4: ConvergenceError, 5: TransformationError, 6: InvalidField,
        7: ConfigurationError, 8: PmError}


def lib():
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} is missing: run __graft_entry__.build()")
        # libb200pm.so links libb200reg.so through its $ORIGIN rpath
        _LIB = C.CDLL(LIB_PATH)
        _LIB.b200pm_icp_create.restype = C.c_void_p
        _LIB.b200pm_icp_destroy.argtypes = [C.c_void_p]
        _LIB.b200pm_icp_load_yaml.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_int]
        _LIB.b200pm_icp_set_default.argtypes = [C.c_void_p, C.c_char_p, C.c_int]
        _LIB.b200pm_icp_describe.argtypes = [C.c_void_p, C.c_char_p, C.c_int]
        _LIB.b200pm_icp_run.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int64, C.c_void_p, C.c_int, C.c_void_p, C.c_int64,
                                        C.c_void_p, C.c_int, C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int),
                                        C.POINTER(C.c_float), C.c_char_p, C.c_int]
        _LIB.b200pm_list_modules.argtypes = [C.c_int, C.c_char_p, C.c_int]
        _LIB.b200pm_module_parameters.argtypes = [C.c_int, C.c_char_p, C.c_char_p, C.c_int, C.c_char_p, C.c_int]
        _LIB.b200pm_apply_filter.argtypes = [C.c_char_p, C.c_char_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int64, C.POINTER(C.c_int64),
                                             C.POINTER(C.c_int), C.c_char_p, C.c_int, C.c_char_p, C.c_int]
    return _LIB


def _raise(code, err):
    if code != 0:
        raise _EXC.get(code, PmError)(err.value.decode("utf-8", "replace"))


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f32(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float32)


def list_modules(kind):
    buf = C.create_string_buffer(8192)
    lib().b200pm_list_modules(KINDS[kind], buf, len(buf))
    return [s for s in buf.value.decode().split("\n") if s]


def module_parameters(kind, name):
    """{param: (default, min, max)} — the static availableParameters() table of a registered module."""
    buf = C.create_string_buffer(16384)
    err = C.create_string_buffer(1024)
    _raise(lib().b200pm_module_parameters(KINDS[kind], name.encode(), buf, len(buf), err, len(err)), err)
    out = {}
    for line in buf.value.decode().split("\n"):
        if line:
            n, d, mn, mx = line.split("|")
            out[n] = (d, mn, mx)
    return out


class ICP:
    """PointMatcher<float>::ICP (PointMatcher.h:729-848)."""

    def __init__(self, yaml_text=None):
        self.h = lib().b200pm_icp_create()
        if yaml_text is not None:
            self.load_yaml(yaml_text)

    def load_yaml(self, text):
        err = C.create_string_buffer(2048)
        _raise(lib().b200pm_icp_load_yaml(self.h, text.encode(), err, len(err)), err)

    def set_default(self):
        err = C.create_string_buffer(2048)
        _raise(lib().b200pm_icp_set_default(self.h, err, len(err)), err)

    def describe(self):
        buf = C.create_string_buffer(8192)
        lib().b200pm_icp_describe(self.h, buf, len(buf))
        return [tuple(s.split(":")) for s in buf.value.decode().split("\n") if s]

    def __call__(self, reading, reference, T_init=None, reading_normals=None, reference_normals=None, force_generic=False):
        rd, rf = _f32(reading), _f32(reference)
        rdn, rfn = _f32(reading_normals), _f32(reference_normals)
        Ti = None if T_init is None else _f32(T_init).reshape(16)
        T = np.zeros(16, dtype=np.float32)
        it, fast, mx = C.c_int(0), C.c_int(0), C.c_int(0)
        ov = C.c_float(0)
        err = C.create_string_buffer(2048)
        code = lib().b200pm_icp_run(self.h, _p(rf), rf.shape[1], _p(rfn), rf.shape[0], _p(rd), rd.shape[1], _p(rdn), rd.shape[0], _p(Ti),
                                    int(force_generic), _p(T), C.byref(it), C.byref(fast), C.byref(mx), C.byref(ov), err, len(err))
        _raise(code, err)
        return T.reshape(4, 4), dict(iterations=it.value, fast_lane=bool(fast.value), max_iterations_reached=bool(mx.value), overlap=ov.value)

    def close(self):
        if self.h:
            lib().b200pm_icp_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def apply_filter(name, params, feat, normals=None):
    """Instantiate a DataPointsFilter by registrar name and run inPlaceFilter on a 4 x N cloud."""
    feat = _f32(feat).copy()
    n = feat.shape[0]
    nrm = np.zeros((n, 3), dtype=np.float32) if normals is None else _f32(normals).copy()
    kv = "".join(f"{k}={v}\n" for k, v in (params or {}).items()).encode()
    n_out, has_n = C.c_int64(0), C.c_int(0)
    names = C.create_string_buffer(4096)
    err = C.create_string_buffer(2048)
    _raise(lib().b200pm_apply_filter(name.encode(), kv, _p(feat), _p(nrm), int(normals is not None), n, C.byref(n_out), C.byref(has_n), names,
                                     len(names), err, len(err)), err)
    m = n_out.value
    labels = [tuple(s.split(":")) for s in names.value.decode().split("\n") if s]
    return feat[:m], (nrm[:m] if has_n.value else None), [(a, int(b)) for a, b in labels]
