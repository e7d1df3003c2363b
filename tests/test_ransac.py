"""Stereo RANSAC rigid alignment (reference: ransac_solver/include/ransac_solver.h:51-166,
ransac_solver/include/stereo_reprojection.h:28-134, caller src/visual_odometer.cpp:93-103).

CPU tests pin the oracle (oracle/ransac_oracle.py) where something can be pinned: the element draws against glibc's own
rand(), the DLT triangulation against the closed form of rectified stereo, the Umeyama fit against known motions and
against a float32 SVD formulation, and Solver::Solve's selection rules on constructed draws.  GPU tests compare
sba_ransac_stereo with the oracle through the C-ABI.  PARITY UNPINNED for cv::triangulatePoints / Eigen::umeyama
(OpenCV and Eigen are absent from the reference tree and from this image)."""
import ctypes

import numpy as np
import pytest

from oracle import ransac_oracle as RO
from stereo_orb_slam_b200 import synthetic
from stereo_orb_slam_b200.ransac import CalcTransformation, GlibcRand, RansacOption, StereoRansac, triangulate_stereo

F = np.float32


# ------------------------------------------------------------------------------------------------ CPU: the oracle, pinned
@pytest.mark.parametrize("seed", [1, 0, 2, 12345, 2 ** 31 + 5, 2 ** 32 - 1])
def test_glibc_rand_restatement_matches_libc(seed):
    """The reference draws `rand() % size` (ransac_solver.h:87); the restated generator is checked against the C
    library this process runs on."""
    libc = ctypes.CDLL("libc.so.6")
    libc.srand(ctypes.c_uint(seed))
    ref = [libc.rand() for _ in range(3000)]
    g = GlibcRand(seed)
    assert [g.rand() for _ in range(3000)] == ref


def test_draws_are_three_per_iteration_in_call_order():
    libc = ctypes.CDLL("libc.so.6")
    libc.srand(1)
    ref = np.array([libc.rand() % 137 for _ in range(300)]).reshape(100, 3)
    assert np.array_equal(GlibcRand(1).draws(100, 137), ref)


def test_oracle_triangulation_recovers_rectified_points():
    """For noise-free rectified observations the DLT null vector is the exact point: x = (u-cx) z / f, z = f b / d."""
    rng = np.random.default_rng(0)
    cam = synthetic.KITTI
    p_l, p_r = synthetic.projection_matrices(cam)
    n = 200
    z = rng.uniform(3, 80, n)
    X = np.stack([rng.uniform(-10, 10, n), rng.uniform(-2, 2, n), z], 1)
    ul = cam["fx"] * X[:, 0] / z + cam["cx"]; vl = cam["fy"] * X[:, 1] / z + cam["cy"]
    ur = cam["fx"] * (X[:, 0] - cam["b"]) / z + cam["cx"]
    got = RO.triangulate(p_l, p_r, np.stack([ul, vl], 1), np.stack([ur, vl], 1))
    # pixels are rounded to float: the depth error is z^2 / (f b) * 2^-24 * u
    tol = (z ** 2 / (cam["fx"] * cam["b"]) * 1.2e-4 + 1e-4)[:, None]
    assert np.all(np.abs(got - X) < tol)


def test_oracle_triangulation_with_row_disparity_minimises_the_algebraic_error():
    """v_l != v_r: no closed form; the result must be the minimiser of |A X| over unit X (checked against eigh)."""
    rng = np.random.default_rng(1)
    p_l, p_r = synthetic.projection_matrices()
    P_l = p_l.reshape(3, 4).astype(F).astype(np.float64); P_r = p_r.reshape(3, 4).astype(F).astype(np.float64)
    o = np.array([[650.2, 190.1, 630.7, 191.0], [100.5, 300.25, 60.0, 299.0]], F)
    got = RO.triangulate(p_l, p_r, o[:, :2], o[:, 2:])
    for k in range(2):
        u_l, v_l, u_r, v_r = o[k].astype(np.float64)
        A = np.stack([u_l * P_l[2] - P_l[0], v_l * P_l[2] - P_l[1], u_r * P_r[2] - P_r[0], v_r * P_r[2] - P_r[1]])
        w, V = np.linalg.eigh(A.T @ A)
        x = V[:, 0] / V[3, 0]
        assert np.allclose(got[k], x[:3], rtol=2e-6)


def test_oracle_umeyama_recovers_rigid_motion_and_is_proper():
    rng = np.random.default_rng(2)
    for n in (3, 4, 50):
        R = synthetic._rodrigues(rng.normal(0, 0.4, (1, 3)))[0]
        t = rng.normal(0, 2, 3)
        src = rng.normal(0, 5, (n, 3)).astype(F)
        dst = (src.astype(np.float64) @ R.T + t).astype(F)
        T = RO.umeyama(src, dst)
        assert np.allclose(T[:3, :3], R, atol=2e-6) and np.allclose(T[:3, 3], t, atol=2e-5)
        assert abs(np.linalg.det(T[:3, :3].astype(np.float64)) - 1.0) < 1e-6
        assert np.array_equal(T[3], [0, 0, 0, 1])
    # a mirrored target still yields a rotation (Eigen's S(2) = -1 branch)
    src = rng.normal(0, 5, (30, 3)).astype(F)
    dst = src * np.array([1, 1, -1], F)
    T = RO.umeyama(src, dst)
    assert abs(np.linalg.det(T[:3, :3].astype(np.float64)) - 1.0) < 1e-6


def test_oracle_double_fit_agrees_with_a_float32_svd_formulation():
    """The reference runs Eigen::umeyama on MatrixXf; the oracle (and the GPU) solve the 3x3 problem in double."""
    pairs, _, _, P_l, P_r = synthetic.make_stereo_pairs(300, 5)
    src = RO.triangulate(P_l, P_r, pairs[:, 0, :2], pairs[:, 0, 2:])
    dst = RO.triangulate(P_l, P_r, pairs[:, 1, :2], pairs[:, 1, 2:])
    S = GlibcRand(1).draws(60, 300)
    worst = 0.0
    for idx in S:
        if len(set(idx.tolist())) < 3:
            continue
        worst = max(worst, float(np.abs(RO.umeyama(src[idx], dst[idx]) - RO.umeyama_float_svd(src[idx], dst[idx])).max()))
    assert worst < 1e-3
    a = RO.umeyama(src, dst); b = RO.umeyama_float_svd(src, dst)
    assert np.abs(a - b).max() < 1e-4


def test_oracle_selection_rules():
    pairs, T_true, bad, P_l, P_r = synthetic.make_stereo_pairs(200, 7)
    a, b = pairs[:, 0], pairs[:, 1]
    good = np.flatnonzero(~bad)
    rng = np.random.default_rng(0)
    S = rng.integers(0, 200, (100, 3))
    S[3] = good[[0, 5, 9]]; S[7] = S[3]; S[40] = S[3]              # the same model three times
    T, labels, losses, info = RO.solve(a, b, P_l, P_r, S)
    c = info["counts"]
    # the FIRST of equal best counts wins (ransac_solver.h:112: count > best_count)
    assert c[3] == c[7] == c[40]
    assert info["best_iteration"] == int(np.argmax(c)) and info["best_count"] == c.max()
    if c[3] == c.max():
        assert info["best_iteration"] <= 3
    assert np.array_equal(labels, losses < F(3.0)) and int(labels.sum()) == info["best_count"]
    # early termination: nothing after the first hypothesis above ceil(size * ratio) runs
    opt = RO.Option(early_termination=True, concensus_ratio=float(c.max() - 1) / 200.0 - 1e-3)
    thr = RO.consensus_threshold(200, opt.concensus_ratio)
    T2, _, _, info2 = RO.solve(a, b, P_l, P_r, S, opt)
    first = int(np.flatnonzero(c > thr)[0])
    assert info2["best_iteration"] == first and np.all(info2["counts"][first + 1:] == -1)
    # final fit = Umeyama over the inliers of the best 3-point model (ransac_solver.h:129-151)
    T3, labels3, _, info3 = RO.solve(a, b, P_l, P_r, S, RO.Option(final_model_fitting=True))
    assert np.array_equal(T3, RO.umeyama(info3["src"][labels], info3["dst"][labels]))
    assert np.abs(T3 - T_true).max() < 0.1


def test_oracle_calc_transformation_needs_ten_pairs():
    pairs, _, _, P_l, P_r = synthetic.make_stereo_pairs(9, 1)
    ok, T, m, l = RO.calc_transformation(pairs[:, 0], pairs[:, 1], P_l, P_r, np.zeros((100, 3), int))
    assert ok is False and T is None
    assert RO.consensus_threshold(10, 0.8) == 8 and RO.consensus_threshold(11, 0.8) == 9 and RO.consensus_threshold(500, 0.8) == 400


# ------------------------------------------------------------------------------------------------ GPU: parity through the C-ABI
def _degenerate(S):
    S = np.asarray(S)
    return (S[:, 0] == S[:, 1]) | (S[:, 0] == S[:, 2]) | (S[:, 1] == S[:, 2])


@pytest.mark.gpu
def test_gpu_triangulation_matches_the_oracle():
    rng = np.random.default_rng(11)
    pairs, _, _, P_l, P_r = synthetic.make_stereo_pairs(4000, 11, z_max=90.0)
    o = pairs[:, 0].copy()
    o[:, 3] += rng.normal(0, 0.4, o.shape[0]).astype(F)            # row disparity: the general DLT case
    got = triangulate_stereo(o, P_l, P_r)
    want = RO.triangulate(P_l, P_r, o[:, :2], o[:, 2:])
    same = np.all(got == want, axis=1).mean()
    assert same > 0.995                                            # the null vector is rounded to float: identical unless a tie
    assert np.all(np.abs(got - want) <= 4e-7 * np.abs(want) + 1e-9)


@pytest.mark.gpu
@pytest.mark.parametrize("n,seed", [(10, 0), (57, 1), (500, 2), (2000, 3)])
def test_gpu_solve_matches_the_oracle(n, seed):
    pairs, T_true, bad, P_l, P_r = synthetic.make_stereo_pairs(n, seed)
    solver = StereoRansac(P_l, P_r)
    S = GlibcRand(1).draws(100, n)
    T, labels, losses = solver.Solve(pairs)                        # draws come from the solver's own GlibcRand(1)
    assert np.array_equal(solver.last["samples"], S)
    To, lo, so, info = RO.solve(pairs[:, 0], pairs[:, 1], P_l, P_r, S)
    cg = np.asarray(solver.last["counts"])
    # per-hypothesis inlier counts: identical, repeated picks (rank-deficient fits, stated completion) included
    assert _degenerate(S).any() or n > 100
    assert np.array_equal(cg, info["counts"])
    assert solver.last["best_iteration"] == info["best_iteration"]
    assert np.abs(T - To).max() <= 2e-7 * max(1.0, np.abs(To).max())
    # labels / losses: bit-identical to the oracle's float arithmetic under the model the GPU returned
    src = RO.triangulate(P_l, P_r, pairs[:, 0, :2], pairs[:, 0, 2:])
    d = RO.reprojection_distance(T, P_l, P_r, src, pairs[:, 1])
    assert np.array_equal(losses, d) and np.array_equal(labels, d < F(3.0))
    if np.array_equal(T, To):
        assert np.array_equal(labels, lo) and np.array_equal(losses, so)
    else:
        assert (labels != lo).sum() <= 1


@pytest.mark.gpu
@pytest.mark.parametrize("final_fit,early", [(True, False), (False, True), (True, True)])
def test_gpu_solve_options_match_the_oracle(final_fit, early):
    pairs, T_true, bad, P_l, P_r = synthetic.make_stereo_pairs(400, 21)
    S = GlibcRand(7).draws(100, 400)
    opt = dict(final_model_fitting=final_fit, early_termination=early, concensus_ratio=0.5)
    solver = StereoRansac(P_l, P_r)
    solver.SetOptions(RansacOption(**opt))
    T, labels, losses = solver.Solve(pairs, samples=S)
    To, lo, so, info = RO.solve(pairs[:, 0], pairs[:, 1], P_l, P_r, S, RO.Option(**opt))
    cg = np.asarray(solver.last["counts"])
    assert np.array_equal(cg, info["counts"])                      # -1 after the early exit on both sides
    assert solver.last["best_iteration"] == info["best_iteration"]
    assert np.abs(T - To).max() <= 5e-7 * max(1.0, np.abs(To).max())
    assert (labels != lo).sum() <= 1 and np.abs(losses - so).max() < 1e-3
    if final_fit:
        assert np.abs(T - T_true).max() < 0.15


@pytest.mark.gpu
def test_gpu_batch_equals_single_problem_calls_bit_for_bit():
    rng = np.random.default_rng(5)
    sizes = [int(s) for s in rng.integers(10, 900, 40)] + [0, 10, 1]
    probs = [synthetic.make_stereo_pairs(s, 100 + i) if s > 0 else None for i, s in enumerate(sizes)]
    P_l, P_r = probs[0][3], probs[0][4]
    firsts = [p[0][:, 0] if p else np.zeros((0, 4), F) for p in probs]
    seconds = [p[0][:, 1] if p else np.zeros((0, 4), F) for p in probs]
    samples = [rng.integers(0, max(s, 1), (100, 3)).astype(np.int32) for s in sizes]
    solver = StereoRansac(P_l, P_r)
    T, labels, losses, info = solver.solve_batch(firsts, seconds, samples)
    for i, s in enumerate(sizes):
        if s == 0:
            assert np.all(T[i][:3] == 0) and labels[i].size == 0
            continue
        Ti, li, si, inf = solver.solve_batch([firsts[i]], [seconds[i]], [samples[i]])
        assert np.array_equal(Ti[0], T[i]) and np.array_equal(li[0], labels[i]) and np.array_equal(si[0], losses[i])
        assert np.array_equal(inf["counts"][0], info["counts"][i])
    # a second run of the same batch: bit-identical (integer / fixed-order reductions only)
    T2, labels2, losses2, _ = solver.solve_batch(firsts, seconds, samples)
    assert np.array_equal(T, T2) and all(np.array_equal(a, b) for a, b in zip(losses, losses2))


@pytest.mark.gpu
def test_gpu_generator_stream_continues_across_calls_and_guard():
    pairs, _, _, P_l, P_r = synthetic.make_stereo_pairs(300, 9)
    solver = StereoRansac(P_l, P_r)
    ok, T1, m1, l1 = CalcTransformation(solver, pairs)
    s1 = solver.last["samples"].copy()
    ok2, T2, m2, l2 = CalcTransformation(solver, pairs)
    s2 = solver.last["samples"].copy()
    g = GlibcRand(1)
    assert ok and ok2 and np.array_equal(s1, g.draws(100, 300)) and np.array_equal(s2, g.draws(100, 300))
    # fewer than 10 pairs: failure, and rand() is not consumed (src/visual_odometer.cpp:100)
    ok3, T3, _, _ = CalcTransformation(solver, pairs[:9])
    assert ok3 is False and T3 is None
    CalcTransformation(solver, pairs)
    assert np.array_equal(solver.last["samples"], g.draws(100, 300))
    # the recovered motion is the frame's relative pose and the mask rejects the mismatches
    _, T_true, bad, _, _ = synthetic.make_stereo_pairs(300, 9)
    assert (m1 == ~bad).mean() > 0.8


@pytest.mark.gpu
def test_gpu_built_in_draw_is_reproducible_and_bad_samples_are_rejected():
    from stereo_orb_slam_b200 import SbaError
    pairs, _, _, P_l, P_r = synthetic.make_stereo_pairs(250, 4)
    solver = StereoRansac(P_l, P_r)
    a = solver.solve_batch([pairs[:, 0]], [pairs[:, 1]], seed=42)
    b = solver.solve_batch([pairs[:, 0]], [pairs[:, 1]], seed=42)
    c = solver.solve_batch([pairs[:, 0]], [pairs[:, 1]], seed=43)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[3]["counts"][0], b[3]["counts"][0])
    assert not np.array_equal(a[3]["counts"][0], c[3]["counts"][0])
    S = np.zeros((100, 3), np.int32); S[50, 1] = 250
    with pytest.raises(SbaError):
        solver.solve_batch([pairs[:, 0]], [pairs[:, 1]], [S])
