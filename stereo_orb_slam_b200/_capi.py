"""ctypes binding of libsba.so (include/sba.h).  No torch types cross this boundary.

The library is the product: if it is missing or a symbol is absent this module raises — there is
no CPU fallback and nothing here imports oracle/.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("SBA_LIB_PATH") or os.path.join(_HERE, "lib", "libsba.so")      # the override is a tuning aid (build variants)

# every symbol include/sba.h declares (tests check the header against this list and the .so)
SYMBOLS = [
    "sba_default_options", "sba_status_string", "sba_last_error", "sba_create", "sba_destroy",
    "sba_set_structure", "sba_set_point_shard", "sba_set_state", "sba_set_measurements", "sba_get_state",
    "sba_run", "sba_solve", "sba_linearize", "sba_schur", "sba_pcg", "sba_time_linearize_schur",
    "sba_get_linearization", "sba_get_schur_pattern", "sba_get_schur_values", "sba_get_sorted_views",
    "sba_num_free_cameras", "sba_launch_count", "sba_comm_unique_id", "sba_comm_init",
    "sba_pack_window", "sba_unpack_window", "sba_matrix_to_pose", "sba_pose_to_matrix",
    "sba_time_stage", "sba_invert_pose", "sba_normalize_pose", "sba_compose_pose", "sba_get_factor_symbolic", "sba_measure_fp64_peak",
    "sba_map_reset", "sba_map_append_frame", "sba_map_num_frames", "sba_map_num_points", "sba_map_optimize", "sba_map_get_poses",
    "sba_map_set_pose", "sba_map_get_points", "sba_map_set_points", "sba_map_last_window", "sba_pgo_optimize",
    "sba_ransac_default_options", "sba_ransac_stereo", "sba_triangulate_stereo", "sba_ransac_stereo_noalloc", "sba_fused_kernel_variant",
]

SBA_OK, SBA_NO_CONVERGENCE, SBA_NUMERICAL_FAILURE, SBA_INVALID_ARGUMENT, SBA_CUDA_ERROR, SBA_NCCL_ERROR = range(6)
TERMINATION = {0: "running", 1: "parameter_tolerance", 2: "function_tolerance", 3: "gradient_tolerance",
               4: "max_iterations", 5: "max_time", 6: "min_radius", 7: "invalid_steps"}


class Options(C.Structure):
    _fields_ = [
        ("max_iterations", C.c_int32), ("max_time_s", C.c_double), ("huber_delta", C.c_double),
        ("point_lower_bound", C.c_double), ("point_upper_bound", C.c_double),
        ("parameter_tolerance", C.c_double), ("function_tolerance", C.c_double),
        ("gradient_tolerance", C.c_double), ("residual_dim", C.c_int32), ("initial_radius", C.c_double),
        ("max_pcg_iterations", C.c_int32), ("pcg_tolerance", C.c_double), ("verbose", C.c_int32), ("linear_solver", C.c_int32),
    ]


class Summary(C.Structure):
    _fields_ = [
        ("initial_cost", C.c_double), ("final_cost", C.c_double), ("iterations", C.c_int32),
        ("successful_steps", C.c_int32), ("linearizations", C.c_int32), ("pcg_iterations", C.c_int32),
        ("termination", C.c_int32), ("num_windows_converged", C.c_int32), ("time_total_s", C.c_double),
        ("time_structure_s", C.c_double), ("time_h2d_s", C.c_double), ("time_d2h_s", C.c_double),
        ("time_solve_s", C.c_double), ("ms_linearize", C.c_double), ("ms_schur", C.c_double),
        ("ms_pcg", C.c_double), ("ms_update", C.c_double),
    ]

    def as_dict(self):
        d = {f: getattr(self, f) for f, _ in self._fields_}
        d["termination_name"] = TERMINATION.get(d["termination"], str(d["termination"]))
        return d


class PgoSummary(C.Structure):
    _fields_ = [("initial_chi2", C.c_double), ("final_chi2", C.c_double), ("final_lambda", C.c_double), ("iterations", C.c_int32),
                ("trials", C.c_int32), ("termination", C.c_int32), ("time_total_s", C.c_double), ("time_symbolic_s", C.c_double)]

    def as_dict(self):
        return {f: getattr(self, f) for f, _ in self._fields_}


class RansacOptions(C.Structure):
    """sba_ransac_options = RANSAC::Option (ransac_solver/include/ransac_solver.h:13-19) + the model's pixel threshold."""
    _fields_ = [("max_iterations", C.c_int32), ("final_model_fitting", C.c_int32), ("early_termination", C.c_int32),
                ("concensus_ratio", C.c_float), ("inlier_threshold", C.c_float)]


class SbaError(RuntimeError):
    def __init__(self, status, msg):
        super().__init__("libsba: status %d (%s): %s" % (status, msg[0], msg[1]))
        self.status = status


_lib = None


def lib():
    """Load libsba.so (raises if it has not been built: `python -c 'import __graft_entry__ as g; g.build()'`)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError("libsba.so not built (%s); run __graft_entry__.build() — there is no fallback" % LIB_PATH)
        L = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
        L.sba_status_string.restype = C.c_char_p
        L.sba_last_error.restype = C.c_char_p
        L.sba_get_schur_pattern.restype = C.c_int64
        L.sba_launch_count.restype = C.c_int64
        for f in ("sba_map_num_frames", "sba_map_num_points", "sba_map_last_window"):
            getattr(L, f).restype = C.c_int64
        L.sba_destroy.restype = None
        L.sba_default_options.restype = None
        L.sba_ransac_default_options.restype = None
        L.sba_matrix_to_pose.restype = None
        L.sba_pose_to_matrix.restype = None
        L.sba_normalize_pose.restype = None
        L.sba_compose_pose.restype = None
        L.sba_create.argtypes = [C.c_int, C.POINTER(C.c_void_p)]
        L.sba_destroy.argtypes = [C.c_void_p]
        L.sba_get_schur_pattern.argtypes = [C.c_void_p, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.c_int64]
        L.sba_num_free_cameras.argtypes = [C.c_void_p]
        L.sba_launch_count.argtypes = [C.c_void_p]
        _lib = L
    return _lib


def check(status, ok=(SBA_OK,)):
    if status not in ok:
        L = lib()
        raise SbaError(status, (L.sba_status_string(status).decode(), L.sba_last_error().decode()))
    return status


def default_options(**kw):
    o = Options()
    lib().sba_default_options(C.byref(o))
    for k, v in kw.items():
        if not hasattr(o, k):
            raise TypeError("unknown option " + k)
        setattr(o, k, v)
    return o


def ptr(a, t):
    return None if a is None else a.ctypes.data_as(C.POINTER(t))


def f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def u8(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.uint8)
