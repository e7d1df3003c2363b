"""ORACLE — test infrastructure only (see sba_oracle.cpp header).  ctypes wrapper.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  PARITY UNPINNED (Ceres absent; reference has no tests) — see DESIGN.md.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_build", "libsba_oracle.so")


class OracleOptions(C.Structure):
    _fields_ = [
        ("max_iterations", C.c_int32),
        ("max_time_s", C.c_double),
        ("huber_delta", C.c_double),
        ("point_lower_bound", C.c_double),
        ("point_upper_bound", C.c_double),
        ("parameter_tolerance", C.c_double),
        ("function_tolerance", C.c_double),
        ("gradient_tolerance", C.c_double),
        ("residual_dim", C.c_int32),
        ("initial_radius", C.c_double),
        ("line_search", C.c_int32),
        ("num_threads", C.c_int32),
        ("analytic_jacobian", C.c_int32),
        ("verbose", C.c_int32),
    ]


class OracleSummary(C.Structure):
    _fields_ = [
        ("initial_cost", C.c_double),
        ("final_cost", C.c_double),
        ("iterations", C.c_int32),
        ("successful_steps", C.c_int32),
        ("linearizations", C.c_int32),
        ("termination", C.c_int32),
        ("reason", C.c_int32),
        ("time_total_s", C.c_double),
        ("time_linearize_s", C.c_double),
        ("time_schur_s", C.c_double),
        ("time_solve_s", C.c_double),
        ("final_radius", C.c_double),
        ("final_gradient_max_norm", C.c_double),
    ]

    def as_dict(self):
        return {f: getattr(self, f) for f, _ in self._fields_}


def build(force=False):
    src = os.path.join(_HERE, "sba_oracle.cpp")
    if force or not os.path.exists(_LIB_PATH) or (
        os.path.exists(src) and os.path.getmtime(src) > os.path.getmtime(_LIB_PATH)
    ):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            build()
        L = C.CDLL(_LIB_PATH)
        L.oracle_cost.restype = C.c_double
        L.oracle_time_linearize_schur.restype = C.c_double
        L.oracle_num_procs.restype = C.c_int
        L.oracle_ls_minimize.restype = C.c_double
        _lib = L
    return _lib


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t))


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _u8(a):
    return np.ascontiguousarray(a, dtype=np.uint8)


def default_options(**kw):
    o = OracleOptions()
    lib().oracle_default_options(C.byref(o))
    for k, v in kw.items():
        setattr(o, k, v)
    return o


def residual(cam, pt, uv, p_l, p_r, mode=0):
    """One residual block, no loss.  Returns r[4], Jc[4,6], Jp[4,3]. mode 0 Jet autodiff, 1 analytic."""
    cam, pt, uv, p_l, p_r = map(_f64, (cam, pt, uv, p_l, p_r))
    r = np.zeros(4); Jc = np.zeros((4, 6)); Jp = np.zeros((4, 3))
    d = C.c_double
    lib().oracle_residual(_p(cam, d), _p(pt, d), _p(uv, d), _p(p_l, d), _p(p_r, d), C.c_int(mode),
                          _p(r, d), _p(Jc, d), _p(Jp, d))
    return r, Jc, Jp


def huber(delta, s):
    rho = C.c_double(); sq = C.c_double()
    lib().oracle_huber(C.c_double(delta), C.c_double(s), C.byref(rho), C.byref(sq))
    return rho.value, sq.value


def ls_minimize(x, f, g, has_g, xmin, xmax):
    x, f, g = _f64(x), _f64(f), _f64(g)
    hg = _i32(has_g)
    d = C.c_double
    return lib().oracle_ls_minimize(x.size, _p(x, d), _p(f, d), _p(g, d), _p(hg, C.c_int32), d(xmin), d(xmax))


def linearize(poses, points, obs_cam, obs_pt, obs_uv, p_l, p_r, huber_delta=1.0, rdim=4, analytic=0):
    poses, points, p_l, p_r = map(_f64, (poses, points, p_l, p_r))
    obs_cam, obs_pt, obs_uv = _i32(obs_cam), _i32(obs_pt), _f32(obs_uv)
    Nc, Np, Nobs = poses.size // 6, points.size // 3, obs_cam.size
    out = dict(r=np.zeros((Nobs, 4)), U=np.zeros((Nc, 6, 6)), V=np.zeros((Np, 3, 3)),
               W=np.zeros((Nobs, 6, 3)), gc=np.zeros((Nc, 6)), gp=np.zeros((Np, 3)))
    cost = C.c_double()
    d = C.c_double
    lib().oracle_linearize(Nc, Np, Nobs, _p(poses, d), _p(points, d), _p(obs_cam, C.c_int32),
                           _p(obs_pt, C.c_int32), _p(obs_uv, C.c_float), _p(p_l, d), _p(p_r, d),
                           d(huber_delta), rdim, analytic, _p(out["r"], d), C.byref(cost),
                           _p(out["U"], d), _p(out["V"], d), _p(out["W"], d), _p(out["gc"], d), _p(out["gp"], d))
    out["cost"] = cost.value
    return out


def cost(poses, points, obs_cam, obs_pt, obs_uv, p_l, p_r, huber_delta=1.0, rdim=4):
    poses, points, p_l, p_r = map(_f64, (poses, points, p_l, p_r))
    obs_cam, obs_pt, obs_uv = _i32(obs_cam), _i32(obs_pt), _f32(obs_uv)
    d = C.c_double
    return lib().oracle_cost(poses.size // 6, points.size // 3, obs_cam.size, _p(poses, d), _p(points, d),
                             _p(obs_cam, C.c_int32), _p(obs_pt, C.c_int32), _p(obs_uv, C.c_float),
                             _p(p_l, d), _p(p_r, d), d(huber_delta), rdim)


def schur(poses, points, obs_cam, obs_pt, obs_uv, p_l, p_r, cam_fixed, radius, huber_delta=1.0, rdim=4):
    poses, points, p_l, p_r = map(_f64, (poses, points, p_l, p_r))
    obs_cam, obs_pt, obs_uv, cam_fixed = _i32(obs_cam), _i32(obs_pt), _f32(obs_uv), _u8(cam_fixed)
    Nc = poses.size // 6
    n = 6 * int((cam_fixed == 0).sum())
    S = np.zeros((n, n)); rhs = np.zeros(n)
    d = C.c_double
    rc = lib().oracle_schur(Nc, points.size // 3, obs_cam.size, _p(poses, d), _p(points, d),
                            _p(obs_cam, C.c_int32), _p(obs_pt, C.c_int32), _p(obs_uv, C.c_float),
                            _p(p_l, d), _p(p_r, d), _p(cam_fixed, C.c_uint8), d(huber_delta), rdim, d(radius),
                            _p(S, d), _p(rhs, d))
    if rc != 0:
        raise RuntimeError("oracle_schur: singular point block")
    return S, rhs


def schur_pattern(Nc, Np, obs_cam, obs_pt, cam_fixed):
    obs_cam, obs_pt, cam_fixed = _i32(obs_cam), _i32(obs_pt), _u8(cam_fixed)
    n = lib().oracle_schur_pattern(Nc, Np, obs_cam.size, _p(obs_cam, C.c_int32), _p(obs_pt, C.c_int32),
                                   _p(cam_fixed, C.c_uint8), None, None, 0)
    rows = np.zeros(n, np.int32); cols = np.zeros(n, np.int32)
    lib().oracle_schur_pattern(Nc, Np, obs_cam.size, _p(obs_cam, C.c_int32), _p(obs_pt, C.c_int32),
                               _p(cam_fixed, C.c_uint8), _p(rows, C.c_int32), _p(cols, C.c_int32), n)
    return rows, cols


def solve(poses, points, obs_cam, obs_pt, obs_uv, p_l, p_r, cam_fixed=None, options=None):
    """Ceres-rule LM solve.  Returns (poses, points, summary dict); inputs are not modified."""
    poses = _f64(poses).copy(); points = _f64(points).copy()
    p_l, p_r = _f64(p_l), _f64(p_r)
    obs_cam, obs_pt, obs_uv = _i32(obs_cam), _i32(obs_pt), _f32(obs_uv)
    Nc = poses.size // 6
    if cam_fixed is None:
        cam_fixed = np.zeros(Nc, np.uint8); cam_fixed[0] = 1
    cam_fixed = _u8(cam_fixed)
    if options is None:
        options = default_options()
    s = OracleSummary()
    d = C.c_double
    rc = lib().oracle_solve(Nc, points.size // 3, obs_cam.size, _p(poses, d), _p(points, d),
                            _p(obs_cam, C.c_int32), _p(obs_pt, C.c_int32), _p(obs_uv, C.c_float),
                            _p(p_l, d), _p(p_r, d), _p(cam_fixed, C.c_uint8), C.byref(options), C.byref(s))
    out = s.as_dict(); out["rc"] = rc
    return poses, points, out


def time_linearize_schur(poses, points, obs_cam, obs_pt, obs_uv, p_l, p_r, cam_fixed, num_threads=0,
                         repeats=1, huber_delta=1.0, rdim=4):
    poses, points, p_l, p_r = map(_f64, (poses, points, p_l, p_r))
    obs_cam, obs_pt, obs_uv, cam_fixed = _i32(obs_cam), _i32(obs_pt), _f32(obs_uv), _u8(cam_fixed)
    d = C.c_double
    return lib().oracle_time_linearize_schur(poses.size // 6, points.size // 3, obs_cam.size, _p(poses, d),
                                             _p(points, d), _p(obs_cam, C.c_int32), _p(obs_pt, C.c_int32),
                                             _p(obs_uv, C.c_float), _p(p_l, d), _p(p_r, d),
                                             _p(cam_fixed, C.c_uint8), d(huber_delta), rdim, num_threads, repeats)


def num_procs():
    return lib().oracle_num_procs()
