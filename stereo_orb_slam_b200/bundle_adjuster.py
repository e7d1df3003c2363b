"""BundleAdjuster — the reference's seam (reference: src/bundle_adjuster.h:13-25), over a StereoMap.

    ba = BundleAdjuster(smap)            # src/slam.cpp:94   (holds a reference to the map, like the C++ class)
    ba.Optimize(start, end)              # src/slam.cpp:123,128,156 ; half-open [start, end)

Pack (src/bundle_adjuster.cpp:41-110) and unpack (:121-132) are libsba.so's sba_pack_window /
sba_unpack_window; the solve (:116-117) is sba_solve on the GPU.  The C++ equivalent of this class is
include/bundle_adjuster.h.  The reference's process-global projections
(ReprojectionError::SetLeftProjection / SetRightProjection, src/reprojection_error.h:43-51) travel
in the map (smap.p_l / smap.p_r).
"""
import ctypes as C

import numpy as np

from . import _capi as capi
from ._capi import ptr
from .solver import Solver

D, I32, I64, F, U32 = C.c_double, C.c_int32, C.c_int64, C.c_float, C.c_uint32


def pack_window(smap, start, end):
    """sba_pack_window on a StereoMap.  Returns dict(poses, points, point_ids, obs_cam, obs_pt, obs_uv)."""
    L = capi.lib()
    start, end = int(start), int(end)
    if start < 0 or end < start or end > smap.num_frames:
        raise ValueError("bad window [%d, %d)" % (start, end))
    T = np.ascontiguousarray(smap.T_wc, np.float32)
    mp = np.ascontiguousarray(smap.points, np.float32)
    off = np.ascontiguousarray(smap.frame_off, np.int64)
    pid = np.ascontiguousarray(smap.obs_point_id, np.uint32)
    uv_in = np.ascontiguousarray(smap.obs_uv, np.float32)
    nobs = int(off[end] - off[start])
    nc = end - start
    poses = np.zeros((nc, 6)); points = np.zeros((max(nobs, 1), 3)); ids = np.zeros(max(nobs, 1), np.uint32)
    oc = np.zeros(max(nobs, 1), np.int32); op = np.zeros(max(nobs, 1), np.int32); uv = np.zeros((max(nobs, 1), 4), np.float32)
    n = I32(0)
    capi.check(L.sba_pack_window(ptr(T, F), ptr(mp, F), ptr(off, I64), ptr(pid, U32), ptr(uv_in, F), I32(start),
                                 I32(end), ptr(poses, D), ptr(points, D), ptr(ids, U32), ptr(oc, I32),
                                 ptr(op, I32), ptr(uv, F), C.byref(n)))
    n = n.value
    return dict(poses=poses, points=points[:n].copy(), point_ids=ids[:n].copy(), obs_cam=oc[:nobs].copy(),
                obs_pt=op[:nobs].copy(), obs_uv=uv[:nobs].copy())


def unpack_window(smap, start, end, poses, points, point_ids):
    """sba_unpack_window: writes T_wc rows [start,end) and the window's map points back into smap (in place)."""
    L = capi.lib()
    T = np.ascontiguousarray(smap.T_wc, np.float32)
    mp = np.ascontiguousarray(smap.points, np.float32)
    poses = np.ascontiguousarray(poses, np.float64); points = np.ascontiguousarray(points, np.float64)
    ids = np.ascontiguousarray(point_ids, np.uint32)
    capi.check(L.sba_unpack_window(ptr(T, F), ptr(mp, F), I32(start), I32(end), ptr(poses, D), ptr(points, D),
                                   ptr(ids, U32), I32(ids.size)))
    smap.T_wc = T.reshape(-1, 4, 4); smap.points = mp.reshape(-1, 3)


def matrix_to_pose(T):
    out = np.zeros(6)
    T = np.ascontiguousarray(T, np.float32)
    capi.lib().sba_matrix_to_pose(ptr(T, F), ptr(out, D))
    return out


def pose_to_matrix(pose):
    out = np.zeros((4, 4), np.float32)
    pose = np.ascontiguousarray(pose, np.float64)
    capi.lib().sba_pose_to_matrix(ptr(pose, D), ptr(out, F))
    return out


class BundleAdjuster:
    """Drop-in for the reference class.  `options` defaults to the reference's solver configuration
    (src/bundle_adjuster.cpp:14-36); the reference ignores the solve status and always writes back —
    so does Optimize(), numerical failures included (last accepted state; see last_summary["status"]); only
    hard errors (CUDA / NCCL / invalid input) raise."""

    def __init__(self, smap, device=0, options=None):
        self.map = smap
        self.options = options if options is not None else capi.default_options()
        self.solver = Solver(device)
        self.last_summary = None
        self.last_window = None

    def Optimize(self, start_frame_id, end_frame_id):
        start, end = int(start_frame_id), int(end_frame_id)
        if end <= start:      # reference: UB on an empty window (src/bundle_adjuster.cpp:113); here a no-op
            self.last_summary = None
            return None
        w = pack_window(self.map, start, end)
        poses, points, summary = self.solver.solve(w["poses"], w["points"], w["obs_cam"], w["obs_pt"], w["obs_uv"],
                                                   self.map.p_l, self.map.p_r, options=self.options)
        unpack_window(self.map, start, end, poses, points, w["point_ids"])
        self.last_summary = summary
        self.last_window = dict(w, poses_out=poses, points_out=points)
        return summary

    optimize = Optimize
