import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")
    # build the oracle (checker) and the emulation helper when missing; libsba.so is built by
    # __graft_entry__.build() and travels prebuilt to the GPU box.
    lib = os.path.join(ROOT, "stereo_orb_slam_b200", "lib", "libsba.so")
    orc = os.path.join(ROOT, "oracle", "_build", "libsba_oracle.so")
    if not os.path.exists(lib) or not os.path.exists(orc):
        import __graft_entry__ as g
        g.build()


@pytest.fixture(scope="session")
def kitti():
    from stereo_orb_slam_b200 import synthetic
    return synthetic.projection_matrices()


def make_window(num_frames, num_points, track_len, seed, **kw):
    """Synthetic map packed as one BA window over all its frames (C-ABI pack)."""
    from stereo_orb_slam_b200 import pack_window, synthetic
    smap = synthetic.make_map(num_frames=num_frames, num_points=num_points, track_len=track_len, seed=seed, **kw)
    w = pack_window(smap, 0, num_frames)
    w["p_l"], w["p_r"], w["map"] = smap.p_l, smap.p_r, smap
    fixed = np.zeros(num_frames, np.uint8); fixed[0] = 1
    w["cam_fixed"] = fixed
    return w


def batch_windows(ws):
    """Independent windows concatenated into one batched problem (config 3 shape): dict with the flat arrays plus
    win_cam_off / win_pt_off."""
    cam_off = np.cumsum([0] + [w["poses"].shape[0] for w in ws]).astype(np.int32)
    pt_off = np.cumsum([0] + [w["points"].shape[0] for w in ws]).astype(np.int32)
    return dict(poses=np.concatenate([w["poses"] for w in ws]), points=np.concatenate([w["points"] for w in ws]),
                obs_cam=np.concatenate([w["obs_cam"] + cam_off[i] for i, w in enumerate(ws)]).astype(np.int32),
                obs_pt=np.concatenate([w["obs_pt"] + pt_off[i] for i, w in enumerate(ws)]).astype(np.int32),
                obs_uv=np.concatenate([w["obs_uv"] for w in ws]), win_cam_off=cam_off, win_pt_off=pt_off,
                p_l=ws[0]["p_l"], p_r=ws[0]["p_r"])


def long_track_map(num_frames=260, num_long=10, num_short=500, seed=5):
    """A map in which `num_long` points are seen by EVERY frame (tracks longer than a chunk's 255-observation cap —
    what loop-closure re-targeting produces, src/camera_frame.h:92-109) plus short-track points: a slowly drifting
    rig in front of a point cloud."""
    from stereo_orb_slam_b200 import synthetic
    rng = np.random.default_rng(seed)
    cam = synthetic.KITTI
    T = np.zeros((num_frames, 4, 4)); T[:, 3, 3] = 1
    f = np.arange(num_frames)
    T[:, :3, :3] = synthetic._rot(0.002 * np.sin(0.05 * f), 0.004 * np.sin(0.03 * f), np.zeros(num_frames))
    T[:, :3, 3] = np.stack([0.4 * np.sin(0.04 * f), 0.05 * np.cos(0.06 * f), 0.01 * f], axis=1)
    R_cw = np.transpose(T[:, :3, :3], (0, 2, 1)); t_cw = -np.einsum("nij,nj->ni", R_cw, T[:, :3, 3])
    n_pts = num_long + num_short
    X = np.stack([rng.uniform(-3, 3, n_pts), rng.uniform(-1, 1, n_pts), rng.uniform(8, 20, n_pts)], axis=1)
    first = np.concatenate([np.zeros(num_long, np.int64), np.sort(rng.integers(0, num_frames - 5, num_short))])
    last = np.concatenate([np.full(num_long, num_frames), first[num_long:] + 5])
    fr, pid = [], []
    for fi in range(num_frames):
        vis = np.nonzero((first <= fi) & (fi < last))[0]
        fr.append(np.full(vis.size, fi)); pid.append(vis)
    fr = np.concatenate(fr); pid = np.concatenate(pid)
    uv, z = synthetic._project(R_cw[fr], t_cw[fr], X[pid], cam)
    uv32 = (uv + rng.normal(0, 0.3, uv.shape)).astype(np.float32)
    frame_off = np.zeros(num_frames + 1, np.int64); np.add.at(frame_off, fr + 1, 1); frame_off = np.cumsum(frame_off)
    T0 = T.copy()
    T0[1:, :3, 3] += rng.normal(0, 0.01, (num_frames - 1, 3))
    p_l, p_r = synthetic.projection_matrices(cam)
    return synthetic.StereoMap(T_wc=T0.astype(np.float32), points=(X + rng.normal(0, 0.05, X.shape)).astype(np.float32),
                               frame_off=frame_off, obs_point_id=pid.astype(np.uint32), obs_uv=uv32,
                               obs_sigma=np.ones(fr.size, np.float32), p_l=p_l, p_r=p_r, gt_T_wc=T, gt_points=X)


@pytest.fixture(scope="session")
def small_window():
    return make_window(8, 600, 5, seed=7)


@pytest.fixture(scope="session")
def emu():
    """tests/emu: the device formulas of csrc/sba_math.cuh compiled for the host (test-only)."""
    import ctypes as C
    out = os.path.join(ROOT, "tests", "emu", "_build", "libemu.so")
    src = os.path.join(ROOT, "tests", "emu", "emu_math.cpp")
    if not os.path.exists(out) or os.path.getmtime(out) < max(os.path.getmtime(src), *[os.path.getmtime(os.path.join(ROOT, 'stereo_orb_slam_b200', 'csrc', f)) for f in ('sba_math.cuh', 'sba_linesearch.cuh')]):
        os.makedirs(os.path.dirname(out), exist_ok=True)
        subprocess.check_call(["/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++", "-O2", "-std=c++17",
                               "-fPIC", "-shared", "-Wno-unknown-pragmas", "-o", out, src])
    L = C.CDLL(out)
    L.emu_obs_cost.restype = C.c_double
    L.emu_ls_minimize.restype = C.c_double
    return L
