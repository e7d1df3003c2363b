"""Synthetic KITTI-like stereo maps in the reference's data model.

A map is what the reference's VisualOdometer::Update leaves behind (reference:
src/visual_odometer.cpp:335-444): a list of frames with a world-from-camera pose T_wc (f32 4x4,
src/camera_frame.h:122-124), a list of map points (f32 xyz, src/map_point.h:43) whose ids are dense
in order of first observation, and per-frame lists of stereo observations
(point_id, u_l, v_l, u_r, v_r, sigma; src/observation.h:11-16).  Here the per-frame lists are stored
flat, frame-major, with `frame_off` giving each frame's slice.
"""
from dataclasses import dataclass, field

import numpy as np

# rectified KITTI-00-like stereo rig; values pass through f32 as the reference's
# projection matrices do (reference: src/camera_model.cpp:38-49, src/slam.cpp:176-209).
KITTI = dict(fx=718.856, fy=718.856, cx=607.1928, cy=185.2157, b=0.53716, width=1241.0, height=376.0)


def projection_matrices(cam=KITTI):
    """p_l, p_r as 12 doubles, row-major 3x4, computed in f32 then widened (reference R6)."""
    fx, fy, cx, cy, b = (np.float32(cam[k]) for k in ("fx", "fy", "cx", "cy", "b"))
    K = np.array([[fx, 0, cx, 0], [0, fy, cy, 0], [0, 0, 1, 0]], np.float32)
    T_r_inv = np.eye(4, dtype=np.float32)
    T_r_inv[0, 3] = -b  # right camera sits at +b along x; P = K [I|0] T_cam^-1
    p_l = K.astype(np.float64).reshape(-1)
    p_r = (K @ T_r_inv).astype(np.float32).astype(np.float64).reshape(-1)
    return p_l, p_r


@dataclass
class StereoMap:
    T_wc: np.ndarray            # [N,4,4] f32 world-from-camera (Frame::GlobalPose)
    points: np.ndarray          # [M,3]   f32 (MapPoint::Position)
    frame_off: np.ndarray       # [N+1]   i64 slice of each frame's observations
    obs_point_id: np.ndarray    # [K]     u32
    obs_uv: np.ndarray          # [K,4]   f32 (u_l, v_l, u_r, v_r)
    obs_sigma: np.ndarray       # [K]     f32 (never read by BA)
    p_l: np.ndarray             # [12] f64
    p_r: np.ndarray             # [12] f64
    gt_T_wc: np.ndarray = None  # ground truth (f64) for known-answer cases
    gt_points: np.ndarray = None
    meta: dict = field(default_factory=dict)

    @property
    def num_frames(self):
        return self.T_wc.shape[0]

    @property
    def num_observations(self):
        return self.obs_point_id.shape[0]


def _rot(rx, ry, rz):
    """R = Ry(yaw) Rx(pitch) Rz(roll) batch; camera convention z forward, x right, y down."""
    cx, sx, cy, sy, cz, sz = np.cos(rx), np.sin(rx), np.cos(ry), np.sin(ry), np.cos(rz), np.sin(rz)
    n = rx.shape[0]
    Rx = np.zeros((n, 3, 3)); Ry = np.zeros((n, 3, 3)); Rz = np.zeros((n, 3, 3))
    Rx[:, 0, 0] = 1; Rx[:, 1, 1] = cx; Rx[:, 1, 2] = -sx; Rx[:, 2, 1] = sx; Rx[:, 2, 2] = cx
    Ry[:, 1, 1] = 1; Ry[:, 0, 0] = cy; Ry[:, 0, 2] = sy; Ry[:, 2, 0] = -sy; Ry[:, 2, 2] = cy
    Rz[:, 2, 2] = 1; Rz[:, 0, 0] = cz; Rz[:, 0, 1] = -sz; Rz[:, 1, 0] = sz; Rz[:, 1, 1] = cz
    return Ry @ Rx @ Rz


def _rodrigues(w):
    th = np.linalg.norm(w, axis=-1)[:, None, None]
    th = np.maximum(th, 1e-12)
    K = np.zeros(w.shape[:-1] + (3, 3))
    K[:, 0, 1] = -w[:, 2]; K[:, 0, 2] = w[:, 1]; K[:, 1, 0] = w[:, 2]
    K[:, 1, 2] = -w[:, 0]; K[:, 2, 0] = -w[:, 1]; K[:, 2, 1] = w[:, 0]
    return np.eye(3) + np.sin(th) / th * K + (1 - np.cos(th)) / th ** 2 * (K @ K)


def trajectory(n, rng, loop=False):
    """Ground-truth T_wc [n,4,4] f64: 0.8-1.2 m/frame forward, yaw random walk 0.5 deg/frame."""
    T = np.zeros((n, 4, 4)); T[:, 3, 3] = 1
    if loop:
        # closed circuit (KITTI-00-like loop): heading turns 2*pi over the sequence, so the last
        # frames re-observe the area of the first ones.
        step = rng.uniform(0.8, 1.2, n)
        yaw = np.linspace(0, 2 * np.pi, n, endpoint=False) + np.cumsum(rng.normal(0, np.deg2rad(0.05), n))
    else:
        step = rng.uniform(0.8, 1.2, n)
        yaw = np.cumsum(rng.normal(0, np.deg2rad(0.5), n)); yaw -= yaw[0]
    pitch = rng.normal(0, np.deg2rad(0.1), n); roll = rng.normal(0, np.deg2rad(0.1), n)
    pitch[0] = roll[0] = 0.0
    R = _rot(pitch, yaw, roll)
    fwd = R[:, :, 2] * step[:, None]
    pos = np.concatenate([np.zeros((1, 3)), np.cumsum(fwd[:-1], axis=0)], axis=0)
    T[:, :3, :3] = R; T[:, :3, 3] = pos
    return T


def _project(T_cw_R, T_cw_t, X, cam):
    """Exact rectified projection of world points X [k,3] by per-row camera (R [k,3,3], t [k,3])."""
    p = np.einsum("kij,kj->ki", T_cw_R, X) + T_cw_t
    z = p[:, 2]
    zs = np.where(np.abs(z) < 1e-9, 1e-9, z)
    ul = cam["fx"] * p[:, 0] / zs + cam["cx"]
    vl = cam["fy"] * p[:, 1] / zs + cam["cy"]
    ur = cam["fx"] * (p[:, 0] - cam["b"]) / zs + cam["cx"]
    return np.stack([ul, vl, ur, vl], axis=1), z


def make_map(num_frames, num_points, track_len, seed, noise_px=0.5, outlier_frac=0.03,
             pose_sigma_t=0.05, pose_sigma_r=0.005, loop=False, loop_frames=0, cam=KITTI,
             noise_free=False, depth_min=4.0, depth_max=80.0, vo_gates=True):
    """Generate a StereoMap.  `noise_free=True` gives the known-answer twin: exact measurements and a
    perturbed start, so the optimum has cost ~ 0 (up to f32 storage of the pixels) at ground truth.

    `vo_gates` (default): the map only holds what the reference's visual odometer lets through —
      * a stereo keypoint pair enters a frame's container only if it triangulates to a depth in [0, 100] m (rectified
        stereo: src/camera_model.cpp:219) and below fx*b (src/visual_odometer.cpp:137, MaxSensibleDistance :111-121);
        that holds for the observation that creates a point and for every re-observation;
      * a re-observation is a RANSAC inlier: 4-term stereo reprojection distance < 3 px
        (ransac_solver/include/stereo_reprojection.h:96-133), so "outliers" are bounded: `outlier_frac` of the
        re-observations get an extra error of 1.5-2.9 px (4-vector norm), well into Huber's linear range (delta = 1),
        and anything at or beyond 3 px from the true projection is dropped.
    `vo_gates=False` is the stress case the reference never sees: disparity clamped at 1 px (depth up to 386 m),
    gross outliers of +-(10-50) px; weakly constrained points then run to the +-1e4 bound."""
    rng = np.random.default_rng(seed)
    gt = trajectory(num_frames, rng, loop=loop)
    R_wc, t_wc = gt[:, :3, :3], gt[:, :3, 3]
    R_cw = np.transpose(R_wc, (0, 2, 1)); t_cw = -np.einsum("nij,nj->ni", R_cw, t_wc)

    # candidates: oversample, keep the first `num_points` that have >= 1 visible observation
    over = int(num_points * 1.6) + 64
    f0 = rng.integers(-(track_len - 1), num_frames, over)          # nominal first frame of the track
    first = np.clip(f0, 0, num_frames - 1)
    depth = rng.uniform(depth_min, depth_max, over)
    u = rng.uniform(0.0, cam["width"], over); v = rng.uniform(0.0, cam["height"], over)
    Xc = np.stack([(u - cam["cx"]) * depth / cam["fx"], (v - cam["cy"]) * depth / cam["fy"], depth], axis=1)
    Xw = np.einsum("kij,kj->ki", R_wc[first], Xc) + t_wc[first]

    j = np.arange(track_len)
    fr = (f0[:, None] + j[None, :]).reshape(-1)
    pid = np.repeat(np.arange(over), track_len)
    if loop and loop_frames > 0:
        # loop closure: points first seen in the first `loop_frames` frames are re-observed by the
        # last frames of the circuit (off-band blocks in the reduced camera matrix)
        early = np.nonzero(first < loop_frames)[0]
        early = early[rng.random(early.size) < 0.3]
        extra_fr = num_frames - loop_frames + first[early]
        fr = np.concatenate([fr, extra_fr]); pid = np.concatenate([pid, early])
    ok = (fr >= 0) & (fr < num_frames)
    fr, pid = fr[ok], pid[ok]
    uv, z = _project(R_cw[fr], t_cw[fr], Xw[pid], cam)
    vis = (z > 0.5) & (z < cam["fx"] * cam["b"]) & (uv[:, 0] >= 0) & (uv[:, 0] < cam["width"]) & \
          (uv[:, 2] >= 0) & (uv[:, 2] < cam["width"]) & (uv[:, 1] >= 0) & (uv[:, 1] < cam["height"])
    fr, pid, uv = fr[vis], pid[vis], uv[vis]
    first_fr = np.full(over, num_frames, np.int64)
    np.minimum.at(first_fr, pid, fr)
    # the measurement noise of every candidate's FIRST observation is drawn up front: that observation creates the
    # point (reference 3.3), and whether the point exists at all depends on the depth it triangulates to
    first_noise = np.zeros((over, 4)) if noise_free else rng.normal(0, noise_px, (over, 4))
    if vo_gates:
        is_first = fr == first_fr[pid]
        m0 = (uv[is_first] + first_noise[pid[is_first]]).astype(np.float32).astype(np.float64)
        disp = m0[:, 0] - m0[:, 2]
        z0 = np.where(disp > 0, cam["fx"] * cam["b"] / np.where(disp > 0, disp, 1.0), np.inf)
        created = np.zeros(over, bool)
        created[pid[is_first]] = (z0 >= 0.0) & (z0 <= 100.0) & (z0 < cam["fx"] * cam["b"])
        keep0 = created[pid]
        fr, pid, uv = fr[keep0], pid[keep0], uv[keep0]
    # dense ids in order of first observation (frame, then candidate id), as VO::Update assigns them
    seen = np.unique(pid)
    if seen.size > num_points:   # uniform subsample (keeps every frame populated), then first-appearance order
        seen = np.sort(rng.choice(seen, num_points, replace=False))
    order = seen[np.lexsort((seen, first_fr[seen]))]
    new_id = np.full(over, -1, np.int64); new_id[order] = np.arange(order.size)
    keep = new_id[pid] >= 0
    fr, pid, uv = fr[keep], new_id[pid[keep]], uv[keep]
    Xw = Xw[order]
    first_noise = first_noise[order]
    # frame-major, point id ascending inside a frame
    o = np.lexsort((pid, fr))
    fr, pid, uv = fr[o], pid[o], uv[o]
    K = fr.size
    if not noise_free:
        noise = rng.normal(0, noise_px, (K, 4))
        out = rng.random(K) < outlier_frac
        # a point's first observation creates it (reference §3.3: triangulated at creation), so it is
        # never an outlier; later re-observations can be.
        fo = np.full(Xw.shape[0], K, np.int64); np.minimum.at(fo, pid, np.arange(K)); out[fo] = False
        noise[fo] = first_noise
        nout = int(out.sum())
        if vo_gates:
            # bounded outliers (RANSAC inliers by construction): error vector of norm 1.5 .. 2.9 px
            dirs = rng.normal(0, 1, (nout, 4)); dirs /= np.linalg.norm(dirs, axis=1, keepdims=True)
            noise[out] = dirs * rng.uniform(1.5, 2.9, (nout, 1))
        else:
            noise[out] += rng.uniform(10, 50, (nout, 4)) * rng.choice([-1.0, 1.0], (nout, 4))
        uv = uv + noise
    else:
        noise = np.zeros((K, 4))
    uv32 = uv.astype(np.float32)
    if vo_gates:
        # every observation is a stereo keypoint pair of the frame's container, which only holds pairs that triangulate
        # inside the range (src/visual_odometer.cpp:137-143 after src/camera_model.cpp:219), and a re-observation is
        # a RANSAC inlier (< 3 px, stereo_reprojection.h:96-133): anything else never reaches the map
        dsp = uv32[:, 0].astype(np.float64) - uv32[:, 2].astype(np.float64)
        okz = (dsp > 0) & (cam["fx"] * cam["b"] / np.where(dsp > 0, dsp, 1.0) <= 100.0)
        okz &= (np.linalg.norm(noise, axis=1) < 3.0)
        okz[fo if not noise_free else np.zeros(0, np.int64)] |= (dsp[fo] > 0) if not noise_free else False
        fr, pid, uv, uv32 = fr[okz], pid[okz], uv[okz], uv32[okz]
        K = fr.size
    frame_off = np.zeros(num_frames + 1, np.int64)
    np.add.at(frame_off, fr + 1, 1); frame_off = np.cumsum(frame_off)

    # initial state: perturbed poses (first one exact), points triangulated from the first noisy
    # observation through the perturbed pose (how the reference creates them, §3.3), stored f32
    dw = rng.normal(0, pose_sigma_r, (num_frames, 3)); dt = rng.normal(0, pose_sigma_t, (num_frames, 3))
    dw[0] = 0; dt[0] = 0
    T0 = gt.copy()
    T0[:, :3, :3] = R_wc @ _rodrigues(dw); T0[:, :3, 3] = t_wc + dt
    first_obs = np.full(Xw.shape[0], K, np.int64)
    np.minimum.at(first_obs, pid, np.arange(K))
    m = uv32[first_obs].astype(np.float64); ff = fr[first_obs]
    d = np.maximum(m[:, 0] - m[:, 2], 1.0)
    zc = cam["fx"] * cam["b"] / d
    Xc0 = np.stack([(m[:, 0] - cam["cx"]) * zc / cam["fx"], (m[:, 1] - cam["cy"]) * zc / cam["fy"], zc], axis=1)
    X0 = np.einsum("kij,kj->ki", T0[ff, :3, :3], Xc0) + T0[ff, :3, 3]
    p_l, p_r = projection_matrices(cam)
    return StereoMap(T_wc=T0.astype(np.float32), points=X0.astype(np.float32), frame_off=frame_off,
                     obs_point_id=pid.astype(np.uint32), obs_uv=uv32, obs_sigma=np.ones(K, np.float32),
                     p_l=p_l, p_r=p_r, gt_T_wc=gt, gt_points=Xw,
                     meta=dict(num_frames=num_frames, num_points=int(Xw.shape[0]), num_observations=int(K),
                               track_len=track_len, seed=int(seed), noise_free=bool(noise_free)))


def make_stereo_pairs(n, seed, outlier_frac=0.3, noise_px=0.1, cam=KITTI, motion=None, z_max=30.0):
    """Temporal matches between two stereo frames as VisualOdometer::Track builds them (src/visual_odometer.cpp:157-171):
    obs_pairs [n,2,4] f32 with [:,0] = the current frame's (u_l, v_l, u_r, v_r) and [:,1] = the previous frame's, so the
    model RANSAC estimates maps current-frame points into the previous frame (= the frame's relative pose).
    Points 4 m to z_max ahead (3-point fits on triangulated points degrade quickly with depth), pixel noise on all four coordinates,
    `outlier_frac` of the pairs mismatched (second observation replaced by another point's).
    Returns (obs_pairs, T_prev_from_curr [4,4] f64, is_outlier [n] bool, P_l [3,4] f32, P_r [3,4] f32)."""
    rng = np.random.default_rng(seed)
    if motion is None:
        w = rng.normal(0, np.deg2rad(1.0), 3)
        t = np.array([rng.normal(0, 0.05), rng.normal(0, 0.02), rng.uniform(0.6, 1.3)])
    else:
        w, t = np.asarray(motion[0], np.float64), np.asarray(motion[1], np.float64)
    T = np.eye(4)
    T[:3, :3] = _rodrigues(w[None])[0]; T[:3, 3] = t
    z = rng.uniform(4.0, z_max, n)
    u = rng.uniform(40, cam["width"] - 40, n); v = rng.uniform(20, cam["height"] - 20, n)
    Xc = np.stack([(u - cam["cx"]) * z / cam["fx"], (v - cam["cy"]) * z / cam["fy"], z], axis=1)     # current camera frame
    Xp = Xc @ T[:3, :3].T + T[:3, 3]                                                                # previous camera frame
    I = np.broadcast_to(np.eye(3), (n, 3, 3)); O = np.zeros((n, 3))
    oc, _ = _project(I, O, Xc, cam)
    op, _ = _project(I, O, Xp, cam)
    oc = oc + rng.normal(0, noise_px, oc.shape); op = op + rng.normal(0, noise_px, op.shape)
    bad = rng.random(n) < outlier_frac
    perm = rng.permutation(n)
    op[bad] = op[perm][bad] + rng.normal(0, 5.0, (int(bad.sum()), 4))
    p_l, p_r = projection_matrices(cam)
    pairs = np.stack([oc, op], axis=1).astype(np.float32)
    return pairs, T, bad, p_l.reshape(3, 4).astype(np.float32), p_r.reshape(3, 4).astype(np.float32)


# BASELINE.json configs: name -> generator arguments
CONFIGS = {
    "C1": dict(num_frames=10, num_points=5000, track_len=8, seed=0xBA00 + 0),
    "C2": dict(num_frames=50, num_points=40000, track_len=7, seed=0xBA00 + 1),
    "C3": dict(num_frames=10, num_points=5000, track_len=8, seed=0xBA00 + 2),   # per window: seed + window id
    "C4": dict(num_frames=4541, num_points=1000000, track_len=7, seed=0xBA00 + 3, loop=True, loop_frames=40),
    # 20 000 frames at ~1 m per frame: a straight open chain would leave the reference's +-1e4 m point box
    # (src/params.h:44,47) and half the map would be clamped at iteration 0 (round 1's "C5 does not converge"); the path
    # is one wide circuit (radius ~3.2 km) WITHOUT loop-closure observations — still an open chain as a graph.
    "C5": dict(num_frames=20000, num_points=5000000, track_len=7, seed=0xBA00 + 4, loop=True, loop_frames=0),
}


def make_config(name, scale=1.0, **over):
    kw = dict(CONFIGS[name]); kw.update(over)
    if scale != 1.0:
        kw["num_points"] = max(8, int(kw["num_points"] * scale))
        if name in ("C4", "C5"):
            kw["num_frames"] = max(4, int(kw["num_frames"] * scale))
    return make_map(**kw)
