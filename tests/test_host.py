"""Host logic through the C-ABI, no GPU: library exports, pack / unpack (reference:
src/bundle_adjuster.cpp:41-110,121-132), pose (de)parameterisation (src/math_utils.h:12-41), and the
structure build (sorted views, Schur block pattern — must be bit-exact)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from oracle import oracle, pack_oracle
from stereo_orb_slam_b200 import Solver, _capi, pack_window, unpack_window, matrix_to_pose, pose_to_matrix, synthetic
from conftest import ROOT, make_window


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "sba.h")).read()
    declared = set(re.findall(r"\b(sba_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(_capi.SYMBOLS), declared ^ set(_capi.SYMBOLS)
    L = _capi.lib()
    for s in declared:
        assert hasattr(L, s), s


def test_default_options_are_the_reference_configuration():
    o = _capi.default_options()
    # src/params.h:34,41 ; src/bundle_adjuster.cpp:35-36,100 ; src/params.h:44,47
    assert (o.max_iterations, o.max_time_s, o.huber_delta) == (50, 1.0, 1.0)
    assert (o.point_lower_bound, o.point_upper_bound) == (-1e4, 1e4)
    assert (o.function_tolerance, o.gradient_tolerance, o.parameter_tolerance) == (1e-16, 1e-16, 1e-8)
    assert o.residual_dim == 4 and o.initial_radius == 1e4
    # the tail of the struct (layout check of the ctypes mirror against the C definition)
    assert (o.max_pcg_iterations, o.pcg_tolerance, o.verbose, o.linear_solver) == (500, 1e-12, 0, 0)


def test_no_cpu_fallback_compute_needs_a_device():
    s = Solver(-1)     # host-only handle: structure yes, compute no
    s.set_structure(2, 1, [0, 1], [0, 0])
    with pytest.raises(_capi.SbaError) as e:
        s.set_state(np.zeros((2, 6)), np.zeros((1, 3)))
    assert e.value.status == _capi.SBA_CUDA_ERROR


def test_pack_indices_bit_exact_incl_duplicates_and_loop_closure():
    smap = synthetic.make_map(num_frames=12, num_points=400, track_len=4, seed=3)
    # duplicate (frame, point) pairs (two features matching one previous feature, visual_odometer.cpp:404-415)
    # and a loop-closure style re-targeting to an old id (camera_frame.h:92-109)
    pid = smap.obs_point_id.copy()
    off = smap.frame_off
    pid[off[5] + 1] = pid[off[5]]
    pid[off[11] + 2] = pid[off[0] + 3]
    smap.obs_point_id = pid
    for (a, b) in ((0, 12), (3, 9), (11, 12), (0, 1), (4, 4)):
        w = pack_window(smap, a, b)
        r = pack_oracle.pack(smap, a, b)
        for k in ("point_ids", "obs_cam", "obs_pt", "obs_uv"):
            assert np.array_equal(w[k], r[k]), (k, a, b)
        assert np.array_equal(w["points"], r["points"].reshape(-1, 3))
        if b > a:
            assert np.abs(w["poses"] - r["poses"]).max() < 5e-6     # f32 arithmetic on both sides
    # first-appearance numbering: the id list is in order of first occurrence in the scan
    w = pack_window(smap, 0, 12)
    seen = []
    for g in smap.obs_point_id:
        if g not in seen[-1:] and g not in set(seen):
            seen.append(int(g))
    assert seen == w["point_ids"].tolist()


def test_pose_round_trip_matches_oracle():
    rng = np.random.default_rng(5)
    for th in (0.0, 1e-6, 1e-3, 0.3, 2.0, 3.1):
        ax = rng.normal(size=3); ax /= np.linalg.norm(ax)
        pose = np.concatenate([ax * th, rng.normal(0, 5, 3)])
        T = pose_to_matrix(pose); T2 = pack_oracle.pose_to_matrix(pose)
        assert np.abs(T - T2).max() < 1e-6
        back = matrix_to_pose(T); back2 = pack_oracle.matrix_to_pose(T2)
        assert np.abs(back - back2).max() < 1e-5
        assert np.abs(back[:3] - pose[:3]).max() < 2e-3 * max(1.0, th) and np.abs(back[3:] - pose[3:]).max() < 1e-6


def test_unpack_writes_window_only():
    smap = synthetic.make_map(num_frames=9, num_points=300, track_len=4, seed=9)
    w = pack_window(smap, 2, 7)
    T0, P0 = smap.T_wc.copy(), smap.points.copy()
    poses = w["poses"] + 1e-3; points = w["points"] + 0.25
    Tr, Pr = pack_oracle.unpack(smap, 2, 7, poses, points, w["point_ids"])
    unpack_window(smap, 2, 7, poses, points, w["point_ids"])
    assert np.abs(smap.T_wc - Tr).max() < 1e-5 and np.array_equal(smap.points, Pr)
    assert np.array_equal(smap.T_wc[:2], T0[:2]) and np.array_equal(smap.T_wc[7:], T0[7:])
    untouched = np.setdiff1d(np.arange(P0.shape[0]), w["point_ids"])
    assert np.array_equal(smap.points[untouched], P0[untouched])
    assert np.allclose(smap.points[w["point_ids"]], (w["points"] + 0.25).astype(np.float32))


@pytest.mark.parametrize("cfg", [dict(num_frames=10, num_points=500, track_len=4, seed=1),
                                 dict(num_frames=30, num_points=800, track_len=6, seed=2, loop=True, loop_frames=5),
                                 # > 2048 free cameras: the pattern comes from the block bitmap, not the dense map
                                 dict(num_frames=2300, num_points=12000, track_len=5, seed=3, loop=True, loop_frames=20)])
def test_structure_sorted_views_and_schur_pattern_bit_exact(cfg):
    w = make_window(**cfg)
    Nc, Np = w["poses"].shape[0], w["points"].shape[0]
    s = Solver(-1).set_structure(Nc, Np, w["obs_cam"], w["obs_pt"])
    bp, bc = s.sorted_views()
    # stable sorts of the reference's scan order by point and by camera
    assert np.array_equal(bp, np.argsort(w["obs_pt"], kind="stable"))
    key = w["obs_cam"].astype(np.int64) * (Np + 1) + w["obs_pt"]
    assert np.array_equal(w["obs_cam"][bc], np.sort(w["obs_cam"]))
    assert np.array_equal(key[bc], np.sort(key, kind="stable"))
    rows, cols = s.schur_pattern()
    r2, c2 = oracle.schur_pattern(Nc, Np, w["obs_cam"], w["obs_pt"], w["cam_fixed"])
    assert np.array_equal(rows, r2) and np.array_equal(cols, c2)
    assert s.num_free_cameras == Nc - 1


def test_structure_rejects_bad_indices_and_handles_degenerate_windows():
    s = Solver(-1)
    with pytest.raises(_capi.SbaError) as e:
        s.set_structure(2, 2, [0, 2], [0, 1])
    assert e.value.status == _capi.SBA_INVALID_ARGUMENT
    with pytest.raises(_capi.SbaError):
        s.set_structure(2, 2, [0, 1], [0, -1])
    # per-frame call (n-1, n): one camera, which is the fixed one => zero free cameras, empty pattern
    s.set_structure(1, 3, [0, 0, 0], [0, 1, 2])
    assert s.num_free_cameras == 0 and s.schur_pattern()[0].size == 0
    # no observations at all
    s.set_structure(3, 0, np.zeros(0, np.int32), np.zeros(0, np.int32))
    assert s.num_free_cameras == 2
    rows, cols = s.schur_pattern()
    assert rows.tolist() == [0, 1] and cols.tolist() == [0, 1]    # diagonal blocks always exist


def test_batched_windows_pattern_is_block_diagonal():
    ws = [make_window(5, 120, 3, seed=20 + i) for i in range(3)]
    cam_off = np.cumsum([0] + [w["poses"].shape[0] for w in ws]).astype(np.int32)
    pt_off = np.cumsum([0] + [w["points"].shape[0] for w in ws]).astype(np.int32)
    oc = np.concatenate([w["obs_cam"] + cam_off[i] for i, w in enumerate(ws)])
    op = np.concatenate([w["obs_pt"] + pt_off[i] for i, w in enumerate(ws)])
    s = Solver(-1).set_structure(cam_off[-1], pt_off[-1], oc, op, win_cam_off=cam_off, win_pt_off=pt_off)
    assert s.num_free_cameras == cam_off[-1] - 3          # first camera of EVERY window is fixed (:113)
    rows, cols = s.schur_pattern()
    fwin = np.repeat(np.arange(3), 4)
    assert np.array_equal(fwin[rows], fwin[cols])
    with pytest.raises(_capi.SbaError):   # an observation may not link two windows
        Solver(-1).set_structure(cam_off[-1], pt_off[-1], oc, op[::-1].copy(), win_cam_off=cam_off, win_pt_off=pt_off)


def test_structure_is_independent_of_the_host_thread_count(monkeypatch):
    """The counting sort into point order (reference scan order kept inside a point, src/bundle_adjuster.cpp:77-102) is
    must not depend on how the structure build is split over the host threads: same tables, and the earliest offending
    observation decides the error message."""
    smap = synthetic.make_config("C2")
    w = pack_window(smap, 0, smap.num_frames)
    Nc, Np = w["poses"].shape[0], w["points"].shape[0]
    out = {}
    for t in (1, 5):
        monkeypatch.setenv("SBA_HOST_THREADS", str(t))
        s = Solver(-1).set_structure(Nc, Np, w["obs_cam"], w["obs_pt"])
        out[t] = s.sorted_views() + s.schur_pattern()
    for a, b in zip(out[1], out[5]):
        assert np.array_equal(a, b)
    assert np.array_equal(out[5][0], np.argsort(w["obs_pt"], kind="stable"))
    oc, op = w["obs_cam"].copy(), w["obs_pt"].copy()
    oc[150000] = Nc + 7; op[100000] = -5          # the earlier observation decides the message
    with pytest.raises(Exception, match="obs_pt out of range"):
        Solver(-1).set_structure(Nc, Np, oc, op)


@pytest.mark.parametrize("k,ok", [(240, True), (241, True), (255, True), (256, True), (700, True), (4097, False)])
def test_long_tracks_are_chunks_of_their_own(k, ok):
    """One point seen by k cameras (loop-closure re-targeting produces long tracks, src/camera_frame.h:92-109; the
    reference has no limit).  Up to 255 observations the point rides in a byte-indexed chunk (round 1 looped forever
    for 241..255); above that it becomes a long-track chunk (one slot per observation, k_*_long kernels); beyond 4096
    the structure build refuses with SBA_INVALID_ARGUMENT instead of crashing."""
    cam = np.concatenate([np.arange(k), np.arange(k)]).astype(np.int32)
    pt = np.concatenate([np.zeros(k), np.ones(k)]).astype(np.int32)          # two such points
    extra_c = np.array([0, 1, 2], np.int32); extra_p = np.array([2, 2, 2], np.int32)
    s = Solver(-1)
    if ok:
        s.set_structure(k, 3, np.concatenate([cam, extra_c]), np.concatenate([pt, extra_p]))
        bp, _ = s.sorted_views()
        assert sorted(bp.tolist()) == list(range(2 * k + 3))
        rows, cols = s.schur_pattern()
        assert rows.size == (k - 1) ** 2                                     # all free cameras co-visible
    else:
        with pytest.raises(_capi.SbaError) as e:
            s.set_structure(k, 3, np.concatenate([cam, extra_c]), np.concatenate([pt, extra_p]))
        assert e.value.status == _capi.SBA_INVALID_ARGUMENT


def test_long_track_pattern_equals_the_oracle_pattern():
    from conftest import long_track_map
    smap = long_track_map(num_frames=270, num_long=3, num_short=60)
    w = pack_window(smap, 0, smap.num_frames)
    s = Solver(-1).set_structure(w["poses"].shape[0], w["points"].shape[0], w["obs_cam"], w["obs_pt"])
    rows, cols = s.schur_pattern()
    fixed = np.zeros(smap.num_frames, np.uint8); fixed[0] = 1
    r2, c2 = oracle.schur_pattern(w["poses"].shape[0], w["points"].shape[0], w["obs_cam"], w["obs_pt"], fixed)
    assert np.array_equal(rows, r2) and np.array_equal(cols, c2)
