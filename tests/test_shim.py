"""The C++ drop-in class (include/bundle_adjuster.h) driven like src/slam.cpp drives the reference's
BundleAdjuster, over test-only Frame/MapPoint stand-ins (tests/shim).  CPU: the pack phase is bit-exact
with the numpy restatement of src/bundle_adjuster.cpp:41-110.  GPU: Optimize() equals the Python host path
(same C-ABI) and writes back with the reference's setter semantics."""
import os
import struct
import subprocess

import numpy as np
import pytest

from stereo_orb_slam_b200 import BundleAdjuster, default_options, synthetic
from oracle import pack_oracle

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DRIVER = os.path.join(ROOT, "tests", "shim", "_build", "shim_driver")


def build_driver():
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "tests", "shim")])
    return DRIVER


def write_map(path, smap):
    with open(path, "wb") as f:
        f.write(struct.pack("<ii", smap.num_frames, smap.points.shape[0]))
        f.write(np.ascontiguousarray(smap.T_wc, np.float32).tobytes())
        f.write(np.ascontiguousarray(smap.frame_off, np.int64).tobytes())
        f.write(np.ascontiguousarray(smap.obs_point_id, np.uint32).tobytes())
        f.write(np.ascontiguousarray(smap.obs_uv, np.float32).tobytes())
        f.write(np.ascontiguousarray(smap.points, np.float32).tobytes())
        f.write(np.ascontiguousarray(smap.p_l, np.float64).tobytes())
        f.write(np.ascontiguousarray(smap.p_r, np.float64).tobytes())


def run_driver(tmp_path, smap, start, end, mode):
    inp, out = str(tmp_path / "map.bin"), str(tmp_path / ("out_%s.bin" % mode))
    write_map(inp, smap)
    r = subprocess.run([build_driver(), inp, out, str(start), str(end), mode], capture_output=True, text=True, timeout=300)
    return r, out


def test_shim_pack_is_bit_exact_with_the_reference_scan_order(tmp_path):
    smap = synthetic.make_map(num_frames=9, num_points=400, track_len=5, seed=23, loop=True, loop_frames=2)
    start, end = 2, 8
    r, out = run_driver(tmp_path, smap, start, end, "pack")
    assert r.returncode == 0, r.stderr
    buf = open(out, "rb").read()
    nc, np_, nobs = struct.unpack_from("<iiq", buf, 0)
    o = 16
    def take(dtype, n):
        nonlocal o
        a = np.frombuffer(buf, dtype, n, o); o += a.nbytes
        return a
    poses = take(np.float64, nc * 6).reshape(nc, 6); points = take(np.float64, np_ * 3).reshape(np_, 3)
    ids = take(np.uint32, np_); cam = take(np.int32, nobs); pt = take(np.int32, nobs); uv = take(np.float32, nobs * 4).reshape(nobs, 4)
    ref = pack_oracle.pack(smap, start, end)
    assert nc == end - start and np_ == ref["points"].shape[0] and nobs == ref["obs_cam"].size
    assert np.array_equal(ids, ref["point_ids"]) and np.array_equal(cam, ref["obs_cam"]) and np.array_equal(pt, ref["obs_pt"])
    assert np.array_equal(uv, ref["obs_uv"]) and np.array_equal(points, ref["points"])
    assert np.abs(poses - ref["poses"]).max() < 2e-6       # f32 angle-axis extraction (src/math_utils.h:12-25)


def test_shim_rejects_a_window_past_the_last_frame(tmp_path):
    smap = synthetic.make_map(num_frames=4, num_points=50, track_len=3, seed=3)
    r, _ = run_driver(tmp_path, smap, 0, 9, "pack")
    assert r.returncode == 3 and "bad window" in r.stderr


@pytest.mark.gpu
def test_shim_optimize_equals_the_python_host_path(tmp_path):
    smap = synthetic.make_map(num_frames=10, num_points=500, track_len=5, seed=19)
    T0, P0 = smap.T_wc.copy(), smap.points.copy()
    start, end = 2, 9
    r, out = run_driver(tmp_path, smap, start, end, "optimize")
    assert r.returncode == 0, r.stderr
    buf = open(out, "rb").read()
    status, final_cost, iters = struct.unpack_from("<idi", buf, 0)
    N, M = smap.num_frames, smap.points.shape[0]
    T = np.frombuffer(buf, np.float32, N * 16, 16).reshape(N, 4, 4)
    P = np.frombuffer(buf, np.float32, M * 3, 16 + N * 64).reshape(M, 3)
    ba = BundleAdjuster(smap, options=default_options(max_time_s=0.0))
    s = ba.Optimize(start, end)
    assert status in (0, 1) and iters == s["iterations"]
    assert abs(final_cost - s["final_cost"]) <= 1e-12 * s["final_cost"]
    ids = ba.last_window["point_ids"]
    assert np.array_equal(P[ids], smap.points[ids])                      # every window point: the BA result in f32
    untouched = np.setdiff1d(np.arange(M), ids)
    assert np.array_equal(P[untouched], P0[untouched])
    assert np.array_equal(T[:start], T0[:start]) and np.array_equal(T[end:], T0[end:])   # half-open window
    assert np.abs(T[start:end] - smap.T_wc[start:end]).max() < 1e-6
