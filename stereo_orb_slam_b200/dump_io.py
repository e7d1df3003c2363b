"""The reference's own problem dump, read and written: lets a map dumped by the reference's
`VisualOdometer::Dump` (src/visual_odometer.cpp:446-505) drive the GPU solver with no glue, and gives the tests a
fixture format a maintainer of the reference already knows.

  poses.txt        N, then per frame the 16 row-major floats of T_wc (Frame::GlobalPose)            :448-462
  points.txt       M, then per map point `x y z` (MapPoint::Position)                                :465-477
  constraints.txt  K, then per observation `frame_id point_id u_l v_l u_r v_r sigma`, frames ascending,
                   observations in stored order                                                       :480-503
  camera file      `key=value` lines (type, fx_l fy_l cx_l cy_l, fx_r ..., b, d_l, d_r, T_l, T_r)    src/camera_utils.cpp:8-64

Values are printed with C++ ostream defaults (6 significant digits), so a dump round trip is lossy in exactly the
way the reference's is; `save_dump(..., precision=9)` keeps float32 exactly.  The projections follow
PinholeCamera (src/camera_model.cpp:38-49): P = K[I|0] * T_cam^-1 in float32, widened to double like
InitializeStereoReprojectionError (src/slam.cpp:176-209).
"""
import os

import numpy as np

from .synthetic import StereoMap


def _fmt(x, precision):
    return ("%." + str(precision) + "g") % float(x)


def save_dump(smap, folder, precision=6):
    """Write poses.txt / points.txt / constraints.txt into `folder` (the reference concatenates folder + name, so a
    trailing separator is implied here)."""
    os.makedirs(folder, exist_ok=True)
    T = np.asarray(smap.T_wc, np.float32).reshape(-1, 16)
    with open(os.path.join(folder, "poses.txt"), "w") as f:
        f.write("%d\n" % T.shape[0])
        for row in T:
            f.write(" ".join(_fmt(v, precision) for v in row) + "\n")
    P = np.asarray(smap.points, np.float32).reshape(-1, 3)
    with open(os.path.join(folder, "points.txt"), "w") as f:
        f.write("%d\n" % P.shape[0])
        for row in P:
            f.write(" ".join(_fmt(v, precision) for v in row) + "\n")
    off = np.asarray(smap.frame_off, np.int64)
    uv = np.asarray(smap.obs_uv, np.float32).reshape(-1, 4)
    pid = np.asarray(smap.obs_point_id)
    sig = np.asarray(smap.obs_sigma, np.float32) if smap.obs_sigma is not None else np.ones(pid.size, np.float32)
    with open(os.path.join(folder, "constraints.txt"), "w") as f:
        f.write("%d\n" % pid.size)
        for i in range(off.size - 1):
            for k in range(int(off[i]), int(off[i + 1])):
                f.write("%d %d %s %s\n" % (i, int(pid[k]), " ".join(_fmt(v, precision) for v in uv[k]), _fmt(sig[k], precision)))


def _numbers(path):
    with open(path) as f:
        return f.read().split()


def load_camera(path):
    """Parse the reference's camera file and return (p_l, p_r) as 12 doubles each, row-major 3x4."""
    vals = {}
    with open(path) as f:
        for line in f:
            line = line.strip()
            if not line or "=" not in line:
                continue
            k, v = line.split("=", 1)
            vals[k.strip()] = v.strip()

    def K(side):
        fx, fy, cx, cy = (np.float32(float(vals[k + "_" + side])) for k in ("fx", "fy", "cx", "cy"))
        return np.array([[fx, 0, cx, 0], [0, fy, cy, 0], [0, 0, 1, 0]], np.float32)

    def T(side):
        if "T_" + side in vals:
            return np.array([float(x) for x in vals["T_" + side].split()], np.float32).reshape(4, 4)
        M = np.eye(4, dtype=np.float32)
        if side == "r" and "b" in vals:
            M[0, 3] = np.float32(float(vals["b"]))     # rectified pair: right camera b to the right of the left one
        return M

    out = []
    for side in ("l", "r"):
        P = (K(side) @ np.linalg.inv(T(side)).astype(np.float32)).astype(np.float32)     # src/camera_model.cpp:49
        out.append(P.astype(np.float64).reshape(12))
    return out[0], out[1]


def load_dump(folder, p_l=None, p_r=None, camera_file=None):
    """Read a dump into a StereoMap.  Projections come from `camera_file` (reference format) or are passed in."""
    if camera_file is not None:
        p_l, p_r = load_camera(camera_file)
    if p_l is None or p_r is None:
        raise ValueError("load_dump needs the projections: pass camera_file or p_l / p_r")
    t = _numbers(os.path.join(folder, "poses.txt"))
    n = int(t[0])
    if len(t) != 1 + 16 * n:
        raise ValueError("poses.txt: expected %d numbers after the count, found %d" % (16 * n, len(t) - 1))
    T = np.array(t[1:], np.float32).reshape(n, 4, 4)
    t = _numbers(os.path.join(folder, "points.txt"))
    m = int(t[0])
    if len(t) != 1 + 3 * m:
        raise ValueError("points.txt: expected %d numbers after the count, found %d" % (3 * m, len(t) - 1))
    P = np.array(t[1:], np.float32).reshape(m, 3)
    t = _numbers(os.path.join(folder, "constraints.txt"))
    k = int(t[0])
    if len(t) != 1 + 7 * k:
        raise ValueError("constraints.txt: expected %d numbers after the count, found %d" % (7 * k, len(t) - 1))
    rows = np.array(t[1:], np.float64).reshape(k, 7)
    fid = rows[:, 0].astype(np.int64)
    pid = rows[:, 1].astype(np.int64)
    if k and (np.any(np.diff(fid) < 0) or fid.min() < 0 or fid.max() >= n):
        raise ValueError("constraints.txt: frame ids must be ascending and < number of poses")
    if k and (pid.min() < 0 or pid.max() >= m):
        raise ValueError("constraints.txt: point id out of range")
    off = np.zeros(n + 1, np.int64)
    np.add.at(off, fid + 1, 1)
    off = np.cumsum(off)
    return StereoMap(T_wc=T, points=P, frame_off=off, obs_point_id=pid.astype(np.uint32), obs_uv=rows[:, 2:6].astype(np.float32),
                     obs_sigma=rows[:, 6].astype(np.float32), p_l=np.asarray(p_l, np.float64), p_r=np.asarray(p_r, np.float64))
