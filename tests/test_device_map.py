"""Device-resident map: the reference's per-frame + sliding-window call pattern (src/slam.cpp:122-129)
replayed on a map that lives inside the handle must leave, after EVERY call, exactly the bits the fresh-pack path
(sba_pack_window + sba_solve + sba_unpack_window on a host map) leaves."""
import numpy as np
import pytest

from stereo_orb_slam_b200 import BundleAdjuster, ResidentBundleAdjuster, Solver, _capi, default_options, replay, synthetic
from stereo_orb_slam_b200.device_map import frame_slices


def test_frame_slices_follow_first_appearance_ids():
    smap = synthetic.make_map(num_frames=12, num_points=500, track_len=5, seed=23)
    tot = 0
    for f, (a, b, first_new, new) in enumerate(frame_slices(smap)):
        assert first_new == tot
        ids = smap.obs_point_id[a:b]
        assert (ids < first_new + new).all()
        tot += new
    assert tot == smap.points.shape[0]


def test_map_calls_need_a_device_and_valid_ids():
    smap = synthetic.make_map(num_frames=4, num_points=60, track_len=3, seed=29)
    L = _capi.lib()
    s = Solver(-1)
    import ctypes as C
    T = np.eye(4, dtype=np.float32)
    rc = L.sba_map_append_frame(s._h, T.ctypes.data_as(C.POINTER(C.c_float)), C.c_int64(0), None, None, C.c_int32(0), None)
    assert rc == _capi.SBA_CUDA_ERROR
    assert L.sba_map_num_frames(s._h) == 0


@pytest.mark.gpu
def test_replay_matches_fresh_pack_solves_bit_for_bit():
    smap = synthetic.make_map(num_frames=25, num_points=2500, track_len=6, seed=37)
    opt = default_options(max_time_s=0.0, max_iterations=12)
    ref = synthetic.make_map(num_frames=25, num_points=2500, track_len=6, seed=37)      # host map driven through the fresh-pack path
    host = BundleAdjuster(ref, options=opt)
    seen = []

    def check(ba, s, e, summ):
        hs = host.Optimize(s, e)
        assert hs["iterations"] == summ["iterations"] and hs["final_cost"] == summ["final_cost"], (s, e)
        assert np.array_equal(ba.poses(s, e - s), ref.T_wc[s:e]), (s, e)
        ids = ba.last_point_ids()
        assert np.array_equal(ids, host.last_window["point_ids"])
        X = ba.points()
        assert np.array_equal(X[ids], ref.points[ids]), (s, e)
        seen.append((s, e))

    ba, calls = replay(smap, interval=10, options=opt, on_call=check)
    assert seen == [(s, e) for s, e, _ in calls]
    assert (0, 10) in seen and (0, 20) in seen and (24, 25) in seen           # sliding windows of src/slam.cpp:126-129
    # untouched parts of the map are exactly what was appended
    assert np.array_equal(ba.points()[ba.num_points - 1], ref.points[-1])
    assert ba.num_frames == 25 and ba.num_points == smap.points.shape[0]
    # the sliding-window call touched free cameras; the per-frame calls did not move any pose
    assert not np.array_equal(ba.poses(1, 9), smap.T_wc[1:10])


@pytest.mark.gpu
def test_resident_map_accepts_pose_and_point_overrides():
    smap = synthetic.make_map(num_frames=6, num_points=300, track_len=4, seed=41)
    ba = ResidentBundleAdjuster(smap.p_l, smap.p_r, options=default_options(max_time_s=0.0))
    for f, (a, b, first_new, new) in enumerate(frame_slices(smap)):
        ba.append_frame(smap.T_wc[f], smap.obs_point_id[a:b], smap.obs_uv[a:b], smap.points[first_new:first_new + new])
    T = smap.T_wc[3].copy(); T[0, 3] += 0.25
    ba.set_pose(3, T)
    assert np.array_equal(ba.poses(3, 1)[0], T)
    X = ba.points(10, 5) + 1.0
    ba.set_points(10, X)
    assert np.array_equal(ba.points(10, 5), X)
    s = ba.Optimize(0, 6)
    assert s["final_cost"] < s["initial_cost"]
