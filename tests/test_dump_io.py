"""The reference's Dump() text format (src/visual_odometer.cpp:446-505) and camera file (src/camera_utils.cpp:8-64)."""
import numpy as np
import pytest

from stereo_orb_slam_b200 import BundleAdjuster, default_options, dump_io, pack_window, synthetic


def test_dump_round_trip_is_exact_at_float32_precision(tmp_path):
    smap = synthetic.make_map(num_frames=6, num_points=120, track_len=4, seed=41)
    dump_io.save_dump(smap, str(tmp_path), precision=9)
    back = dump_io.load_dump(str(tmp_path), p_l=smap.p_l, p_r=smap.p_r)
    assert np.array_equal(back.T_wc, smap.T_wc.astype(np.float32)) and np.array_equal(back.points, smap.points.astype(np.float32))
    assert np.array_equal(back.frame_off, smap.frame_off) and np.array_equal(back.obs_point_id, smap.obs_point_id)
    assert np.array_equal(back.obs_uv, smap.obs_uv.astype(np.float32))
    a, b = pack_window(smap, 1, 5), pack_window(back, 1, 5)
    for k in a:
        assert np.array_equal(a[k], b[k]), k


def test_dump_with_reference_formatting_loses_what_the_reference_loses(tmp_path):
    smap = synthetic.make_map(num_frames=4, num_points=60, track_len=3, seed=42)
    dump_io.save_dump(smap, str(tmp_path))                 # ostream default: 6 significant digits
    lines = open(tmp_path / "constraints.txt").read().splitlines()
    assert int(lines[0]) == smap.num_observations and len(lines[1].split()) == 7
    assert len(open(tmp_path / "poses.txt").read().splitlines()[1].split()) == 16
    back = dump_io.load_dump(str(tmp_path), p_l=smap.p_l, p_r=smap.p_r)
    assert np.abs(back.obs_uv - smap.obs_uv).max() < 1e-2 and np.abs(back.points - smap.points).max() < 1e-3 * np.abs(smap.points).max()
    assert np.array_equal(back.obs_point_id, smap.obs_point_id)


def test_camera_file_gives_the_kitti_projections(tmp_path):
    cam = synthetic.KITTI
    path = tmp_path / "camera.txt"
    path.write_text("type=StereoRectified\nfx_l=%r\nfy_l=%r\ncx_l=%r\ncy_l=%r\nfx_r=%r\nfy_r=%r\ncx_r=%r\ncy_r=%r\nb=%r\n"
                    "d_l=0 0 0 0 0\nd_r=0 0 0 0 0\nT_l=1 0 0 0 0 1 0 0 0 0 1 0 0 0 0 1\nT_r=1 0 0 %r 0 1 0 0 0 0 1 0 0 0 0 1\n" % (
                        cam["fx"], cam["fy"], cam["cx"], cam["cy"], cam["fx"], cam["fy"], cam["cx"], cam["cy"], cam["b"], cam["b"]))
    p_l, p_r = dump_io.load_camera(str(path))
    ref_l, ref_r = synthetic.projection_matrices()
    assert np.allclose(p_l, ref_l, rtol=0, atol=1e-4) and np.allclose(p_r, ref_r, rtol=0, atol=1e-3)
    assert p_r[3] < 0 and abs(p_r[3] + cam["fx"] * cam["b"]) < 1e-3


def test_bad_dumps_are_rejected(tmp_path):
    smap = synthetic.make_map(num_frames=3, num_points=20, track_len=2, seed=43)
    dump_io.save_dump(smap, str(tmp_path), precision=9)
    with pytest.raises(ValueError):
        dump_io.load_dump(str(tmp_path))                   # no projections
    txt = open(tmp_path / "constraints.txt").read().splitlines()
    open(tmp_path / "constraints.txt", "w").write("\n".join(txt[:-1]) + "\n")
    with pytest.raises(ValueError):
        dump_io.load_dump(str(tmp_path), p_l=smap.p_l, p_r=smap.p_r)


@pytest.mark.gpu
def test_a_dumped_map_solves_like_the_original(tmp_path):
    smap = synthetic.make_map(num_frames=8, num_points=400, track_len=5, seed=44)
    dump_io.save_dump(smap, str(tmp_path), precision=9)
    back = dump_io.load_dump(str(tmp_path), p_l=smap.p_l, p_r=smap.p_r)
    opt = default_options(max_time_s=0.0)
    s1 = BundleAdjuster(smap, options=opt).Optimize(0, 8)
    s2 = BundleAdjuster(back, options=opt).Optimize(0, 8)
    assert s1["final_cost"] == s2["final_cost"] and s1["iterations"] == s2["iterations"]
    assert np.array_equal(smap.T_wc, back.T_wc) and np.array_equal(smap.points, back.points)
