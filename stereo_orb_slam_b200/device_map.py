"""ResidentBundleAdjuster — the reference's call pattern on a device-resident map.

    ba = ResidentBundleAdjuster(p_l, p_r)
    ba.append_frame(T_wc, obs_point_id, obs_uv, new_points)     # what VisualOdometer::Update adds per frame
    ba.Optimize(n - 1, n)                                       # src/slam.cpp:123
    if n % interval == 0: ba.Optimize(max(0, n - 2 * interval), n)   # src/slam.cpp:126-129

Map points and stereo pixels stay on the GPU between calls (libsba.so sba_map_*); poses and the integer index are
mirrored on the host inside the library.  `replay(smap, ...)` drives a whole StereoMap through that pattern.
"""
import ctypes as C

import numpy as np

from . import _capi as capi
from ._capi import ptr
from .solver import Solver

D, I32, I64, F, U32 = C.c_double, C.c_int32, C.c_int64, C.c_float, C.c_uint32


class ResidentBundleAdjuster:
    def __init__(self, p_l, p_r, device=0, options=None):
        self.solver = Solver(device)
        self._L, self._h = self.solver._L, self.solver._h
        self.p_l, self.p_r = capi.f64(p_l), capi.f64(p_r)
        self.options = options if options is not None else capi.default_options()
        capi.check(self._L.sba_map_reset(self._h))
        self.last_summary = None

    @property
    def num_frames(self):
        return int(self._L.sba_map_num_frames(self._h))

    @property
    def num_points(self):
        return int(self._L.sba_map_num_points(self._h))

    def append_frame(self, T_wc, obs_point_id, obs_uv, new_points):
        T = np.ascontiguousarray(T_wc, np.float32).reshape(16)
        pid = np.ascontiguousarray(obs_point_id, np.uint32); uv = np.ascontiguousarray(obs_uv, np.float32).reshape(-1, 4)
        newp = np.ascontiguousarray(new_points, np.float32).reshape(-1, 3)
        capi.check(self._L.sba_map_append_frame(self._h, ptr(T, F), I64(pid.size), ptr(pid, U32), ptr(uv, F), I32(newp.shape[0]),
                                                ptr(newp, F)))

    def Optimize(self, start, end):
        s = capi.Summary()
        rc = self._L.sba_map_optimize(self._h, I32(int(start)), I32(int(end)), ptr(self.p_l, D), ptr(self.p_r, D),
                                      C.byref(self.options), C.byref(s))
        capi.check(rc, ok=(capi.SBA_OK, capi.SBA_NO_CONVERGENCE, capi.SBA_NUMERICAL_FAILURE))
        out = s.as_dict(); out["status"] = rc
        ms = (C.c_double * 5)()
        n = int(self._L.sba_map_last_window(self._h, None, I64(0), ms))
        out["num_points"] = n
        out["ms_phases"] = dict(zip(("pack", "structure", "state", "solve", "unpack"), (float(v) for v in ms)))
        self.last_summary = out
        return out

    optimize = Optimize

    def last_point_ids(self):
        n = int(self._L.sba_map_last_window(self._h, None, I64(0), None))
        ids = np.zeros(max(n, 1), np.uint32)
        self._L.sba_map_last_window(self._h, ptr(ids, U32), I64(n), None)
        return ids[:n]

    def poses(self, first=0, count=None):
        count = self.num_frames - first if count is None else count
        T = np.zeros((count, 4, 4), np.float32)
        capi.check(self._L.sba_map_get_poses(self._h, I32(first), I32(count), ptr(T, F)))
        return T

    def set_pose(self, frame, T_wc):
        T = np.ascontiguousarray(T_wc, np.float32).reshape(16)
        capi.check(self._L.sba_map_set_pose(self._h, I32(frame), ptr(T, F)))

    def points(self, first=0, count=None):
        count = self.num_points - first if count is None else count
        X = np.zeros((count, 3), np.float32)
        capi.check(self._L.sba_map_get_points(self._h, I64(first), I64(count), ptr(X, F)))
        return X

    def set_points(self, first, xyz):
        X = np.ascontiguousarray(xyz, np.float32).reshape(-1, 3)
        capi.check(self._L.sba_map_set_points(self._h, I64(first), I64(X.shape[0]), ptr(X, F)))


def frame_slices(smap):
    """Per frame: (observation slice, number of points the frame creates).  Ids are dense in order of first
    observation (src/visual_odometer.cpp:349-372), so frame f creates ids [max id seen so far + 1 ...]."""
    out, seen = [], 0
    for f in range(smap.num_frames):
        a, b = int(smap.frame_off[f]), int(smap.frame_off[f + 1])
        top = int(smap.obs_point_id[a:b].max()) + 1 if b > a else seen
        new = max(top - seen, 0)
        out.append((a, b, seen, new))
        seen = max(seen, top)
    return out


def replay(smap, interval=10, device=0, options=None, on_call=None):
    """The reference's main loop (src/slam.cpp:108-129) on a resident map: per frame append + Optimize(n-1, n), every
    `interval` frames the sliding window Optimize(max(0, n - 2 interval), n).  Returns the adjuster and the list of
    (start, end, summary) of every call."""
    ba = ResidentBundleAdjuster(smap.p_l, smap.p_r, device=device, options=options)
    calls = []
    for f, (a, b, first_new, new) in enumerate(frame_slices(smap)):
        ba.append_frame(smap.T_wc[f], smap.obs_point_id[a:b], smap.obs_uv[a:b], smap.points[first_new:first_new + new])
        n = f + 1
        windows = [(n - 1, n)]
        if n % interval == 0:
            windows.append((max(0, n - 2 * interval), n))
        for (s, e) in windows:
            summ = ba.Optimize(s, e)
            calls.append((s, e, summ))
            if on_call is not None:
                on_call(ba, s, e, summ)
    return ba, calls
