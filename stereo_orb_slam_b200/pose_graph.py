"""Pose-graph optimisation on the GPU — the solve PoseGraphOptimizer::Optimize hands to g2o
(reference: src/pose_graph_optimizer.cpp:34-96).

    T_opt, summary = optimize_pose_graph(T_wc, edges, Z)        # vertex 0 fixed, 10 Levenberg iterations, Huber(1)

`odometry_graph(T_wc, T_rel)` builds what AddOdometryConstraints builds (:98-170): one vertex per frame, one edge
(i-1 -> i) per frame with the frame's relative pose as measurement; loop-closure edges (:172-217) are appended by the
caller as (id_2 -> id_1, measured transform), the vertex order the reference uses (:192-198).
"""
import ctypes as C

import numpy as np

from . import _capi as capi
from ._capi import ptr
from .solver import Solver

D, I32, U8 = C.c_double, C.c_int32, C.c_uint8


def optimize_pose_graph(T_wc, edges, Z, fixed=None, information=None, huber_delta=1.0, iterations=10, device=0, solver=None):
    """T_wc [n,4,4] world-from-camera poses, edges [m,2] (i, j), Z [m,4,4] measured X_i^-1 X_j.
    Returns (optimised poses [n,4,4] f64, summary dict)."""
    T = np.ascontiguousarray(np.asarray(T_wc, np.float64).reshape(-1, 4, 4)).copy()
    E = np.ascontiguousarray(np.asarray(edges, np.int32).reshape(-1, 2))
    Zm = np.ascontiguousarray(np.asarray(Z, np.float64).reshape(-1, 4, 4))
    if Zm.shape[0] != E.shape[0]:
        raise ValueError("one measurement per edge")
    ei = np.ascontiguousarray(E[:, 0]); ej = np.ascontiguousarray(E[:, 1])
    fx = None if fixed is None else np.ascontiguousarray(fixed, np.uint8)
    info = None if information is None else np.ascontiguousarray(information, np.float64).reshape(6)
    sv = solver if solver is not None else Solver(device)
    s = capi.PgoSummary()
    rc = sv._L.sba_pgo_optimize(sv._h, I32(T.shape[0]), ptr(T, D), I32(E.shape[0]), ptr(ei, I32), ptr(ej, I32), ptr(Zm, D),
                                ptr(fx, U8), ptr(info, D), D(huber_delta), I32(iterations), C.byref(s))
    capi.check(rc)
    return T, s.as_dict()


def odometry_graph(T_wc, T_rel=None):
    """Vertices and odometry edges as AddOdometryConstraints builds them (src/pose_graph_optimizer.cpp:98-170).  T_rel[i] is
    frame i's RelativePose (pose in the previous frame); by default it is derived from the global poses."""
    T = np.asarray(T_wc, np.float64).reshape(-1, 4, 4)
    n = T.shape[0]
    edges = np.stack([np.arange(0, n - 1), np.arange(1, n)], axis=1).astype(np.int32)
    if T_rel is None:
        Z = np.stack([np.linalg.inv(T[i]) @ T[i + 1] for i in range(n - 1)]) if n > 1 else np.zeros((0, 4, 4))
    else:
        Z = np.asarray(T_rel, np.float64).reshape(-1, 4, 4)[1:]
    return edges, Z
