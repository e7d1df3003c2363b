"""Parity tests proper: the CUDA path (through the C-ABI) against the CPU oracle on the same seeded inputs.

Tolerances (BASELINE.json north_star): final cost 1e-5 relative, poses 1e-4 m / 1e-5 rad, points 1e-4 m;
indices / sparsity pattern bit-exact.  Stage-level blocks are compared at 1e-9 relative (fp64, different
summation order)."""
import numpy as np
import pytest

from oracle import oracle
from stereo_orb_slam_b200 import BundleAdjuster, Solver, _capi, default_options, pack_window, synthetic
from conftest import batch_windows, long_track_map, make_window

pytestmark = pytest.mark.gpu

PARITY = default_options(max_time_s=0.0)        # the reference's 1 s wall clock is not reproducible (params.h:40)


def oracle_opts(**kw):
    return oracle.default_options(max_time_s=0.0, num_threads=0, **kw)


def load(w, device=0):
    s = Solver(device).set_structure(w["poses"].shape[0], w["points"].shape[0], w["obs_cam"], w["obs_pt"])
    s.set_state(w["poses"], w["points"]); s.set_measurements(w["obs_uv"], w["p_l"], w["p_r"])
    return s


def relerr(a, b):
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-300)


@pytest.mark.parametrize("rdim", [4, 3])
def test_linearization_blocks(small_window, rdim):
    w = small_window
    s = load(w)
    o = default_options(residual_dim=rdim)
    s.linearize(o)
    g = s.get_linearization()
    ref = oracle.linearize(w["poses"], w["points"], w["obs_cam"], w["obs_pt"], w["obs_uv"], w["p_l"], w["p_r"], rdim=rdim)
    for k in ("V", "W", "gp"):
        assert relerr(g[k], ref[k]) < 1e-9, k
    for k in ("U", "gc"):       # camera blocks exist for FREE cameras only (the fixed one is not an unknown, :113)
        assert relerr(g[k][1:], ref[k][1:]) < 1e-9, k
        assert not g[k][0].any()
    assert abs(g["cost"][0] - ref["cost"]) <= 1e-12 * ref["cost"]
    assert np.abs(g["r"] - ref["r"]).max() < 1e-9


def test_kat_residuals_on_device():
    """Golden vectors (mpmath) through the CUDA kernels: one camera + one point + one observation each."""
    import json, os
    kats = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "reprojection_kats.json")))["cases"]
    for c in kats:
        s = Solver(0).set_structure(1, 1, [0], [0], cam_fixed=[0])
        s.set_state(np.array(c["cam"]), np.array(c["point"]))
        s.set_measurements(np.array(c["uv"], np.float32), c["p_l"], c["p_r"])
        s.linearize()
        g = s.get_linearization()
        r = np.array(c["r"]); Jc = np.array(c["Jc"]); Jp = np.array(c["Jp"])
        sq = min(1.0, (r @ r) ** -0.25)        # Huber corrector sqrt(rho')
        assert np.abs(g["r"][0] - sq * r).max() < 1e-10, c["name"]
        for got, ref in ((g["U"][0], sq * sq * Jc.T @ Jc), (g["V"][0], sq * sq * Jp.T @ Jp), (g["W"][0], sq * sq * Jc.T @ Jp),
                         (g["gc"][0], sq * sq * Jc.T @ r), (g["gp"][0], sq * sq * Jp.T @ r)):
            assert relerr(got, ref) < 1e-10, c["name"]


def test_schur_complement_and_pattern(small_window):
    w = small_window
    s = load(w)
    s.linearize()
    # the stage entry applies the damping in the Jacobi-scaled space; compare at a huge radius where damping
    # vanishes (1e-6 floor / radius), against the oracle's dense S at the same radius
    radius = 1e12
    s.schur(radius)
    S, rhs, _ = s.schur_dense()
    So, rhso = oracle.schur(w["poses"], w["points"], w["obs_cam"], w["obs_pt"], w["obs_uv"], w["p_l"], w["p_r"],
                            w["cam_fixed"], radius)
    assert relerr(S, So) < 1e-8 and relerr(rhs, rhso) < 1e-8
    rows, cols = s.schur_pattern()
    r2, c2 = oracle.schur_pattern(w["poses"].shape[0], w["points"].shape[0], w["obs_cam"], w["obs_pt"], w["cam_fixed"])
    assert np.array_equal(rows, r2) and np.array_equal(cols, c2)
    # structural zeros of the oracle's dense S are exactly the blocks outside the pattern
    nf = s.num_free_cameras
    nzb = np.abs(So.reshape(nf, 6, nf, 6)).max(axis=(1, 3)) > 0
    pat = np.zeros((nf, nf), bool); pat[rows, cols] = True
    assert np.array_equal(nzb, pat)


def test_pcg_solves_the_reduced_system(small_window):
    w = small_window
    s = load(w)
    s.linearize(); s.schur(1e4)
    it = s.pcg(default_options(pcg_tolerance=1e-12, max_pcg_iterations=500, linear_solver=1))
    S, rhs, x = s.schur_dense()
    xs = np.linalg.solve(S, rhs)
    assert 0 < it <= 500
    assert np.abs(x - xs).max() <= 1e-8 * np.abs(xs).max()


@pytest.mark.parametrize("path", ["cholesky", "cholesky_window", "cholesky_cluster", "cholesky_sparse", "window", "cluster", "multilaunch", "cooperative"])
def test_every_linear_solver_path_solves_the_same_system(small_window, path, monkeypatch):
    """The reduced system goes through the block Cholesky (windows of <= 64 free cameras, the default there: the
    banded envelope in one CTA's shared memory when it fits, else one CTA per window on an L2 workspace, or one
    8-CTA cluster per window from 16 cameras up; the block-sparse factorisation above 64 cameras) or one of four PCG
    paths depending on its size (one CTA per window, one 8-CTA cluster per window, one cooperative launch over the
    global block-CSR, or the multi-launch path); force each on the same small window."""
    for k in ("SBA_NO_CLUSTER", "SBA_FORCE_CLUSTER", "SBA_FORCE_MULTILAUNCH", "SBA_FORCE_COOP_PCG", "SBA_NO_DIRECT", "SBA_FORCE_SPARSE_CHOL", "SBA_NO_BAND"):
        monkeypatch.delenv(k, raising=False)
    if path == "cholesky_window":       # "cholesky" itself is the banded single-CTA kernel (k_chol_band); this is k_chol_window
        monkeypatch.setenv("SBA_NO_BAND", "1")
    if path == "cholesky_sparse":       # the block-sparse factorisation of the global configs, forced on a small window
        monkeypatch.setenv("SBA_FORCE_SPARSE_CHOL", "1")
    elif path in ("cluster", "cholesky_cluster"):
        monkeypatch.setenv("SBA_FORCE_CLUSTER", "1")
    elif path == "multilaunch":
        monkeypatch.setenv("SBA_FORCE_MULTILAUNCH", "1")
    elif path == "cooperative":
        monkeypatch.setenv("SBA_FORCE_COOP_PCG", "1")
    elif path == "window":
        monkeypatch.setenv("SBA_NO_CLUSTER", "1")
    chol = path.startswith("cholesky")
    solver = 2 if chol else 1
    w = small_window
    s = load(w)
    s.linearize(); s.schur(1e4)
    it = s.pcg(default_options(pcg_tolerance=1e-12, max_pcg_iterations=500, linear_solver=solver))
    S, rhs, x = s.schur_dense()
    xs = np.linalg.solve(S, rhs)
    assert (it == 0) if chol else (0 < it <= 500)
    assert np.abs(x - xs).max() <= (1e-11 if chol else 1e-8) * np.abs(xs).max()
    po, xo, so = oracle.solve(w["poses"], w["points"], w["obs_cam"], w["obs_pt"], w["obs_uv"], w["p_l"], w["p_r"], options=oracle_opts())
    opt = default_options(max_time_s=0.0, linear_solver=solver)
    poses, points, summ = Solver(0).solve(w["poses"], w["points"], w["obs_cam"], w["obs_pt"], w["obs_uv"], w["p_l"], w["p_r"], options=opt)
    check_solution(w, summ, poses, points, so, po, xo)


@pytest.mark.parametrize("frames", [5, 24])
def test_cholesky_of_a_batch_of_windows_equals_pcg(frames, monkeypatch):
    """Config-3 shape: several independent windows in one call, one CTA per window in k_chol_window — with the
    lower block-triangle in shared memory (<= 20 free cameras) or in the L2-resident workspace (more; the cluster
    kernel is switched off to get there)."""
    monkeypatch.setenv("SBA_NO_CLUSTER", "1")
    monkeypatch.setenv("SBA_NO_BAND", "1")
    ws = [make_window(frames + i, 300 + 50 * i, 4, 40 + i) for i in range(4)]
    cam_off = np.cumsum([0] + [x["poses"].shape[0] for x in ws]).astype(np.int32)
    pt_off = np.cumsum([0] + [x["points"].shape[0] for x in ws]).astype(np.int32)
    poses = np.concatenate([x["poses"] for x in ws]); points = np.concatenate([x["points"] for x in ws])
    cam = np.concatenate([x["obs_cam"] + cam_off[i] for i, x in enumerate(ws)]).astype(np.int32)
    pt = np.concatenate([x["obs_pt"] + pt_off[i] for i, x in enumerate(ws)]).astype(np.int32)
    uv = np.concatenate([x["obs_uv"] for x in ws])
    out = {}
    for name, solver in (("chol", 2), ("pcg", 1)):
        s = Solver(0).set_structure(poses.shape[0], points.shape[0], cam, pt, win_cam_off=cam_off, win_pt_off=pt_off)
        s.set_state(poses, points); s.set_measurements(uv, ws[0]["p_l"], ws[0]["p_r"])
        s.linearize(); s.schur(1e4); s.pcg(default_options(pcg_tolerance=1e-13, max_pcg_iterations=500, linear_solver=solver))
        out[name] = s.schur_values()[2].copy()
    assert np.abs(out["chol"] - out["pcg"]).max() <= 1e-8 * np.abs(out["pcg"]).max()


@pytest.mark.parametrize("frames,track", [(5, 4), (24, 4), (50, 7), (30, 30)])
def test_banded_cholesky_of_a_batch_of_windows_equals_numpy(frames, track):
    """k_chol_band: windows of different sizes in one call, narrow bands (track length 4 / 7) and a fully co-visible
    window (track length = frames: the envelope is the whole lower triangle), each against numpy on its own block."""
    ws = [make_window(frames + i, 300 + 50 * i, track, 140 + i) for i in range(3)]
    b = batch_windows(ws)
    s = Solver(0).set_structure(b["poses"].shape[0], b["points"].shape[0], b["obs_cam"], b["obs_pt"],
                                win_cam_off=b["win_cam_off"], win_pt_off=b["win_pt_off"])
    s.set_state(b["poses"], b["points"]); s.set_measurements(b["obs_uv"], b["p_l"], b["p_r"])
    s.linearize(); s.schur(1e4)
    assert s.pcg(default_options(linear_solver=2)) == 0
    S, rhs, x = s.schur_dense()
    xs = np.linalg.solve(S, rhs)
    assert np.abs(x - xs).max() <= 1e-10 * np.abs(xs).max()
    s.pcg(default_options(linear_solver=2))
    assert np.array_equal(s.schur_values()[2], x)                       # bit-reproducible


@pytest.mark.parametrize("case", ["few_slots", "many_slots", "small_chunks"])
def test_fused_chunk_kernel_equals_the_two_kernel_path_bit_for_bit(small_window, monkeypatch, case):
    """k_fused_chunk (K1-K3 of a chunk in one CTA, W kept in registers) against k_lin_chunk + k_schur_chunk (W through
    HBM): same operations in the same order, so the reduced system and the whole LM trajectory agree to the bit.
    few_slots: the chunk's index tables are staged in shared memory; many_slots (every point seen by 30 cameras: 465
    local blocks per chunk): they are read from global memory; small_chunks: the 128-thread variant."""
    monkeypatch.delenv("SBA_NO_FUSED", raising=False)
    monkeypatch.delenv("SBA_OBS_CAP", raising=False)
    if case == "small_chunks":
        monkeypatch.setenv("SBA_OBS_CAP", "100")
    w = make_window(30, 300, 30, seed=17) if case == "many_slots" else small_window
    s = load(w)
    s.linearize(); s.schur(1e4)
    ref = s.schur_values()
    lin_ref = s.get_linearization(residuals=False)
    s.time_stage(8, 1, False, 1e4)                      # one bench step = fused kernel + assembly
    got = s.schur_values()
    lin = s.get_linearization(residuals=False)
    assert np.array_equal(got[0], ref[0]) and np.array_equal(got[1], ref[1])
    for k in ("U", "V", "W", "gc", "gp"):
        assert np.array_equal(lin[k], lin_ref[k]), k
    args = (w["poses"], w["points"], w["obs_cam"], w["obs_pt"], w["obs_uv"], w["p_l"], w["p_r"])
    a = Solver(0).solve(*args, options=PARITY)
    monkeypatch.setenv("SBA_NO_FUSED", "1")
    b = Solver(0).solve(*args, options=PARITY)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]) and a[2]["final_cost"] == b[2]["final_cost"]
    assert a[2]["iterations"] == b[2]["iterations"]


def test_cooperative_pcg_is_bit_reproducible_on_a_global_problem():
    """k_pcg_coop on thousands of block-rows: every CTA sums the per-row dots itself, so a fast CTA must not
    overwrite them for the next phase while a slow one is still reading (the two phases use separate buffers).
    Two solves of the same system must agree to the bit."""
    smap = synthetic.make_config("C4", scale=0.5)
    b = pack_window(smap, 0, smap.num_frames)
    s = Solver(0).set_structure(b["poses"].shape[0], b["points"].shape[0], b["obs_cam"], b["obs_pt"])
    s.set_state(b["poses"], b["points"]); s.set_measurements(b["obs_uv"], smap.p_l, smap.p_r)
    s.linearize(); s.schur(1e4)
    xs = []
    for _ in range(3):
        it = s.pcg(default_options(max_pcg_iterations=200, linear_solver=1))
        xs.append(s.schur_values()[2].copy())
        assert 0 < it <= 200
    assert np.array_equal(xs[0], xs[1]) and np.array_equal(xs[0], xs[2])


def check_solution(w, summary, poses, points, so, po, xo):
    assert abs(summary["final_cost"] - so["final_cost"]) <= 1e-5 * so["final_cost"]
    assert np.abs(poses[:, :3] - po[:, :3]).max() < 1e-5          # rad
    assert np.abs(poses[:, 3:] - po[:, 3:]).max() < 1e-4          # m
    assert np.abs(points - xo).max() < 1e-4                       # m
    assert np.array_equal(poses[0], w["poses"][0])                # gauge: first camera untouched (:113)


def test_full_solve_matches_oracle_small(small_window):
    w = small_window
    poses, points, s = Solver(0).solve(w["poses"], w["points"], w["obs_cam"], w["obs_pt"], w["obs_uv"], w["p_l"], w["p_r"],
                                       options=PARITY)
    po, xo, so = oracle.solve(w["poses"], w["points"], w["obs_cam"], w["obs_pt"], w["obs_uv"], w["p_l"], w["p_r"],
                              options=oracle_opts())
    assert s["status"] == _capi.SBA_OK and so["termination"] == 0
    assert abs(s["initial_cost"] - so["initial_cost"]) <= 1e-12 * so["initial_cost"]
    check_solution(w, s, poses, points, so, po, xo)


def test_full_solve_config1_local_window():
    """BASELINE.json configs[0]: 10 keyframes / 5k points / ~20k observations."""
    smap = synthetic.make_config("C1")
    w = pack_window(smap, 0, smap.num_frames)
    poses, points, s = Solver(0).solve(w["poses"], w["points"], w["obs_cam"], w["obs_pt"], w["obs_uv"], smap.p_l, smap.p_r,
                                       options=PARITY)
    po, xo, so = oracle.solve(w["poses"], w["points"], w["obs_cam"], w["obs_pt"], w["obs_uv"], smap.p_l, smap.p_r,
                              options=oracle_opts())
    check_solution(w, s, poses, points, so, po, xo)


def test_noise_free_recovers_ground_truth():
    w = make_window(6, 300, 4, seed=11, noise_free=True)
    poses, points, s = Solver(0).solve(w["poses"], w["points"], w["obs_cam"], w["obs_pt"], w["obs_uv"], w["p_l"], w["p_r"],
                                       options=PARITY)
    assert s["final_cost"] / w["obs_cam"].size < 1e-6
    gt = w["map"].gt_points[w["point_ids"]]
    assert np.median(np.abs(points - gt)) < 1e-3


def test_three_residual_variant_matches_oracle(small_window):
    w = small_window
    poses, points, s = Solver(0).solve(w["poses"], w["points"], w["obs_cam"], w["obs_pt"], w["obs_uv"], w["p_l"], w["p_r"],
                                       options=default_options(max_time_s=0.0, residual_dim=3))
    po, xo, so = oracle.solve(w["poses"], w["points"], w["obs_cam"], w["obs_pt"], w["obs_uv"], w["p_l"], w["p_r"],
                              options=oracle_opts(residual_dim=3))
    check_solution(w, s, poses, points, so, po, xo)


def test_zero_free_cameras_per_frame_call():
    """ba.Optimize(n-1, n) (src/slam.cpp:123): the only camera is the fixed one => points-only refinement."""
    smap = synthetic.make_map(num_frames=6, num_points=400, track_len=4, seed=13)
    w = pack_window(smap, 5, 6)
    poses, points, s = Solver(0).solve(w["poses"], w["points"], w["obs_cam"], w["obs_pt"], w["obs_uv"], smap.p_l, smap.p_r,
                                       options=PARITY)
    po, xo, so = oracle.solve(w["poses"], w["points"], w["obs_cam"], w["obs_pt"], w["obs_uv"], smap.p_l, smap.p_r,
                              options=oracle_opts())
    assert np.array_equal(poses, w["poses"])
    assert abs(s["final_cost"] - so["final_cost"]) <= 1e-5 * max(so["final_cost"], 1e-12)
    assert np.abs(points - xo).max() < 1e-4


def test_duplicate_pairs_and_single_observation_points():
    # depth <= 25 m keeps every point well constrained, so both solvers converge (a flat valley makes the
    # end-game decisions rounding-noise driven and the 1e-4 m point tolerance meaningless)
    smap = synthetic.make_map(num_frames=7, num_points=300, track_len=4, seed=17, depth_max=25.0)
    pid = smap.obs_point_id.copy(); off = smap.frame_off
    pid[off[3] + 1] = pid[off[3]]; pid[off[4] + 5] = pid[off[4] + 4]
    smap.obs_point_id = pid
    w = pack_window(smap, 0, 7)
    poses, points, s = Solver(0).solve(w["poses"], w["points"], w["obs_cam"], w["obs_pt"], w["obs_uv"], smap.p_l, smap.p_r,
                                       options=PARITY)
    po, xo, so = oracle.solve(w["poses"], w["points"], w["obs_cam"], w["obs_pt"], w["obs_uv"], smap.p_l, smap.p_r,
                              options=oracle_opts())
    check_solution(w, s, poses, points, so, po, xo)


def test_batched_windows_equal_individual_solves():
    """config 3 shape: independent windows in one call, each with its own LM state — bitwise equal to solving
    each window alone (fixed-order reductions, no atomics on the data path)."""
    ws = [make_window(6, 250, 4, seed=30 + i) for i in range(5)]
    cam_off = np.cumsum([0] + [w["poses"].shape[0] for w in ws]).astype(np.int32)
    pt_off = np.cumsum([0] + [w["points"].shape[0] for w in ws]).astype(np.int32)
    oc = np.concatenate([w["obs_cam"] + cam_off[i] for i, w in enumerate(ws)])
    op = np.concatenate([w["obs_pt"] + pt_off[i] for i, w in enumerate(ws)])
    uv = np.concatenate([w["obs_uv"] for w in ws])
    P = np.concatenate([w["poses"] for w in ws]); X = np.concatenate([w["points"] for w in ws])
    sv = Solver(0)
    pb, xb, sb = sv.solve(P, X, oc, op, uv, ws[0]["p_l"], ws[0]["p_r"], win_cam_off=cam_off, win_pt_off=pt_off, options=PARITY)
    tot = 0.0; conv = 0
    for i, w in enumerate(ws):
        p1, x1, s1 = sv.solve(w["poses"], w["points"], w["obs_cam"], w["obs_pt"], w["obs_uv"], w["p_l"], w["p_r"], options=PARITY)
        assert np.array_equal(p1, pb[cam_off[i]:cam_off[i + 1]])
        assert np.array_equal(x1, xb[pt_off[i]:pt_off[i + 1]])
        tot += s1["final_cost"]; conv += s1["num_windows_converged"]
    assert abs(tot - sb["final_cost"]) <= 1e-12 * tot
    assert conv == sb["num_windows_converged"]


def test_run_is_deterministic_and_idempotent(small_window):
    w = small_window
    a = Solver(0).solve(w["poses"], w["points"], w["obs_cam"], w["obs_pt"], w["obs_uv"], w["p_l"], w["p_r"], options=PARITY)
    b = Solver(0).solve(w["poses"], w["points"], w["obs_cam"], w["obs_pt"], w["obs_uv"], w["p_l"], w["p_r"], options=PARITY)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    # idempotence: a converged state is (numerically) a fixed point of a second solve
    c = Solver(0).solve(a[0], a[1], w["obs_cam"], w["obs_pt"], w["obs_uv"], w["p_l"], w["p_r"], options=PARITY)
    assert abs(c[2]["final_cost"] - a[2]["final_cost"]) <= 1e-9 * a[2]["final_cost"]
    assert np.abs(c[1] - a[1]).max() < 1e-4


def test_bundle_adjuster_optimize_writes_back_like_the_reference():
    smap = synthetic.make_map(num_frames=10, num_points=500, track_len=5, seed=19)
    T0, P0 = smap.T_wc.copy(), smap.points.copy()
    ba = BundleAdjuster(smap, options=PARITY)
    s = ba.Optimize(2, 9)                                    # half-open: frame 9 untouched (src/slam.cpp:156)
    assert s["final_cost"] < s["initial_cost"]
    assert np.array_equal(smap.T_wc[9], T0[9]) and np.array_equal(smap.T_wc[:2], T0[:2])
    assert not np.array_equal(smap.T_wc[3:9], T0[3:9])
    assert np.abs(smap.T_wc[2] - T0[2]).max() < 1e-4         # fixed camera: only re-normalised through f32
    ids = ba.last_window["point_ids"]
    assert np.array_equal(smap.points[ids], ba.last_window["points_out"].astype(np.float32))
    untouched = np.setdiff1d(np.arange(P0.shape[0]), ids)
    assert np.array_equal(smap.points[untouched], P0[untouched])
    assert ba.Optimize(4, 4) is None                          # empty window: no-op


def c2_window(**kw):
    smap = synthetic.make_config("C2", **kw)
    w = pack_window(smap, 0, smap.num_frames)
    w["p_l"], w["p_r"] = smap.p_l, smap.p_r
    return w


def test_config2_full_size_matches_oracle_at_termination():
    """BASELINE.json configs[1] at full size (50 keyframes / 40k points / ~200k obs), generated under the reference's
    own gates (points and observations its visual odometer would have dropped do not exist: src/camera_model.cpp:219,
    src/visual_odometer.cpp:137, stereo_reprojection.h:96-133).  GPU and oracle are compared at the solver's own
    termination — the reference's 50-iteration cap (src/params.h:34) or a tolerance, whichever comes first — at the
    north_star tolerances, and the two must stop for the same reason after the same number of iterations."""
    w = c2_window()
    args = (w["poses"], w["points"], w["obs_cam"], w["obs_pt"], w["obs_uv"], w["p_l"], w["p_r"])
    sv = Solver(0)
    pg, xg, sg = sv.solve(*args, options=PARITY)
    po, xo, so = oracle.solve(*args, options=oracle_opts())
    assert sg["iterations"] == so["iterations"] and sg["termination"] == so["reason"]
    assert sg["successful_steps"] == so["successful_steps"]
    check_solution(w, sg, pg, xg, so, po, xo)
    # the reported cost is the oracle's cost function at the GPU's solution
    c = oracle.cost(pg, xg, w["obs_cam"], w["obs_pt"], w["obs_uv"], w["p_l"], w["p_r"])
    assert abs(c - sg["final_cost"]) <= 1e-10 * c
    assert sg["final_cost"] < 0.1 * sg["initial_cost"] and np.abs(xg).max() < 1e3      # nothing runs to the +-1e4 bound


def test_config2_stress_without_the_reference_gates():
    """The ungated variant (far points at 1 px disparity, gross +-10..50 px outliers — inputs the reference never
    produces): weakly constrained points run to the +-1e4 bound (src/bundle_adjuster.cpp:104-108) and the end game is
    decided by rounding, so GPU and oracle are compared at iteration 16, where both still walk the same
    trajectory, and only size-independent properties are asserted at the 50-iteration cap."""
    w = c2_window(vo_gates=False)
    sv = Solver(0)
    args = (w["poses"], w["points"], w["obs_cam"], w["obs_pt"], w["obs_uv"], w["p_l"], w["p_r"])
    p16, x16, s16 = sv.solve(*args, options=default_options(max_time_s=0.0, max_iterations=16))
    po, xo, so = oracle.solve(*args, options=oracle_opts(max_iterations=16))
    assert s16["iterations"] == 16 and so["iterations"] == 16
    check_solution(w, s16, p16, x16, so, po, xo)
    p50, x50, s50 = sv.solve(*args, options=PARITY)
    assert s50["status"] in (_capi.SBA_OK, _capi.SBA_NO_CONVERGENCE) and s50["iterations"] <= 50
    assert s50["final_cost"] <= s16["final_cost"] * (1 + 1e-12) and s50["final_cost"] < 0.5 * s50["initial_cost"]
    c = oracle.cost(p50, x50, w["obs_cam"], w["obs_pt"], w["obs_uv"], w["p_l"], w["p_r"])
    assert abs(c - s50["final_cost"]) <= 1e-10 * c
    assert np.abs(x50).max() <= 1e4                                                    # the bounds hold


def test_run_ahead_lm_loop_equals_the_synchronous_loop_bit_for_bit(monkeypatch):
    """Single GPU: the next linearisation is enqueued before the host reads the step's outcome (a rejected step
    re-linearises at the same point, a finished window ignores it).  The trajectory — rejected steps, line searches,
    windows finishing at different iterations — must equal the synchronous loop's to the bit."""
    cases = []
    w = make_window(8, 600, 5, 3, vo_gates=False, noise_px=2.0)    # rejected steps (47 of 50 succeed), points at the bounds
    cases.append(((w["poses"], w["points"], w["obs_cam"], w["obs_pt"], w["obs_uv"], w["p_l"], w["p_r"]), {}))
    b = batch_windows([make_window(6 + (i % 3), 300 + 40 * i, 4, 70 + i) for i in range(12)])
    cases.append(((b["poses"], b["points"], b["obs_cam"], b["obs_pt"], b["obs_uv"], b["p_l"], b["p_r"]),
                  dict(win_cam_off=b["win_cam_off"], win_pt_off=b["win_pt_off"])))
    for args, kw in cases:
        monkeypatch.delenv("SBA_NO_RUN_AHEAD", raising=False)
        a = Solver(0).solve(*args, options=PARITY, **kw)
        monkeypatch.setenv("SBA_NO_RUN_AHEAD", "1")
        c = Solver(0).solve(*args, options=PARITY, **kw)
        assert np.array_equal(a[0], c[0]) and np.array_equal(a[1], c[1])
        for k in ("final_cost", "iterations", "successful_steps", "linearizations", "termination"):
            assert a[2][k] == c[2][k], k
        assert a[2]["successful_steps"] < a[2]["iterations"] or kw      # the stress window does reject steps


def test_config3_many_windows_match_the_oracle():
    """BASELINE.json configs[2] shape: 256 DISTINCT 10-keyframe windows (seed + window id) in one batched call, every
    window with its own trust region; 16 of them, spread over the batch, are checked against oracle.solve at the
    north_star tolerances, and the batch total against the sum of its windows."""
    nwin = 256
    seed0 = synthetic.CONFIGS["C3"]["seed"]
    ws = []
    for i in range(nwin):
        smap = synthetic.make_config("C3", seed=seed0 + i, num_points=1250)       # quarter-size windows keep the test short
        w = pack_window(smap, 0, smap.num_frames)
        w["p_l"], w["p_r"] = smap.p_l, smap.p_r
        ws.append(w)
    b = batch_windows(ws)
    pb, xb, sb = Solver(0).solve(b["poses"], b["points"], b["obs_cam"], b["obs_pt"], b["obs_uv"], b["p_l"], b["p_r"],
                                 win_cam_off=b["win_cam_off"], win_pt_off=b["win_pt_off"], options=PARITY)
    assert sb["status"] in (_capi.SBA_OK, _capi.SBA_NO_CONVERGENCE)
    co, po_ = b["win_cam_off"], b["win_pt_off"]
    for i in range(0, nwin, nwin // 16):
        w = ws[i]
        po, xo, so = oracle.solve(w["poses"], w["points"], w["obs_cam"], w["obs_pt"], w["obs_uv"], w["p_l"], w["p_r"],
                                  options=oracle_opts())
        pg, xg = pb[co[i]:co[i + 1]], xb[po_[i]:po_[i + 1]]
        c = oracle.cost(pg, xg, w["obs_cam"], w["obs_pt"], w["obs_uv"], w["p_l"], w["p_r"])
        assert abs(c - so["final_cost"]) <= 1e-5 * so["final_cost"], i
        assert np.abs(pg[:, :3] - po[:, :3]).max() < 1e-5 and np.abs(pg[:, 3:] - po[:, 3:]).max() < 1e-4, i
        assert np.abs(xg - xo).max() < 1e-4, i
        assert np.array_equal(pg[0], w["poses"][0])
    tot = sum(oracle.cost(pb[co[i]:co[i + 1]], xb[po_[i]:po_[i + 1]], w["obs_cam"], w["obs_pt"], w["obs_uv"], w["p_l"], w["p_r"])
              for i, w in enumerate(ws))
    assert abs(tot - sb["final_cost"]) <= 1e-10 * tot


def global_case(name, scale, **kw):
    smap = synthetic.make_config(name, scale=scale, **kw)
    w = pack_window(smap, 0, smap.num_frames)
    w["p_l"], w["p_r"] = smap.p_l, smap.p_r
    return w


@pytest.mark.parametrize("name,scale", [("C4", 0.25), ("C5", 0.05)])
def test_global_configs_scaled_match_the_oracle(name, scale):
    """BASELINE.json configs[3] / [4] shapes — C4: 1135 poses / 250k points / 1.4M observations with the loop closed;
    C5: a 1000-pose open chain / 250k points.  The reduced camera system (6 804 / 5 994 unknowns) is factored exactly
    on both sides, as the reference does (SPARSE_SCHUR + CHOLMOD, src/bundle_adjuster.cpp:26-29): block-sparse Cholesky
    in nested-dissection order on the GPU, row-envelope Cholesky in the oracle.  Compared at the solver's own
    termination at the north_star tolerances."""
    w = global_case(name, scale)
    args = (w["poses"], w["points"], w["obs_cam"], w["obs_pt"], w["obs_uv"], w["p_l"], w["p_r"])
    sv = Solver(0)
    pg, xg, sg = sv.solve(*args, options=PARITY)
    assert sg["pcg_iterations"] == 0                              # exact factorisation, no PCG
    po, xo, so = oracle.solve(*args, options=oracle_opts())
    assert sg["iterations"] == so["iterations"] and sg["termination"] == so["reason"]
    check_solution(w, sg, pg, xg, so, po, xo)
    # and the PCG path north_star names reaches the same minimum when it is given the iterations it needs
    pp, xp, sp = sv.solve(*args, options=default_options(max_time_s=0.0, linear_solver=1, max_pcg_iterations=4000, max_iterations=8))
    p8, x8, s8 = sv.solve(*args, options=default_options(max_time_s=0.0, max_iterations=8))
    assert abs(sp["final_cost"] - s8["final_cost"]) <= 1e-6 * s8["final_cost"]


def test_sparse_cholesky_solves_the_reduced_system_of_a_loop():
    """k_spchol against numpy on a 227-pose loop (C4 at 1/20): S x = rhs to 1e-10, and bit-identical on a re-run."""
    w = global_case("C4", 0.05)
    s = load(w)
    s.linearize(); s.schur(1e4)
    assert s.pcg(default_options(linear_solver=2)) == 0
    S, rhs, x = s.schur_dense()
    xs = np.linalg.solve(S, rhs)
    assert np.abs(x - xs).max() <= 1e-10 * np.abs(xs).max()
    s.pcg(default_options(linear_solver=2))
    assert np.array_equal(s.schur_values()[2], x)


@pytest.mark.parametrize("fused", [True, False])
def test_long_tracks_match_the_oracle(fused, monkeypatch):
    """Points seen by 260 frames (> 255 observations: the long-track kernels) next to ordinary short tracks: blocks
    and the full solve against the oracle, through the fused and the two-kernel linearisation."""
    if not fused:
        monkeypatch.setenv("SBA_NO_FUSED", "1")
    smap = long_track_map()
    w = pack_window(smap, 0, smap.num_frames)
    w["p_l"], w["p_r"] = smap.p_l, smap.p_r
    s = load(w)
    s.linearize()
    g = s.get_linearization()
    ref = oracle.linearize(w["poses"], w["points"], w["obs_cam"], w["obs_pt"], w["obs_uv"], w["p_l"], w["p_r"])
    for k in ("V", "W", "gp"):
        assert relerr(g[k], ref[k]) < 1e-9, k
    for k in ("U", "gc"):
        assert relerr(g[k][1:], ref[k][1:]) < 1e-9, k
    assert abs(g["cost"][0] - ref["cost"]) <= 1e-12 * ref["cost"]
    fixed = np.zeros(smap.num_frames, np.uint8); fixed[0] = 1
    s.schur(1e12)
    S, rhs, _ = s.schur_dense()
    So, rhso = oracle.schur(w["poses"], w["points"], w["obs_cam"], w["obs_pt"], w["obs_uv"], w["p_l"], w["p_r"], fixed, 1e12)
    assert relerr(S, So) < 1e-8 and relerr(rhs, rhso) < 1e-8
    args = (w["poses"], w["points"], w["obs_cam"], w["obs_pt"], w["obs_uv"], w["p_l"], w["p_r"])
    pg, xg, sg = Solver(0).solve(*args, options=default_options(max_time_s=0.0, max_iterations=15))
    po, xo, so = oracle.solve(*args, options=oracle_opts(max_iterations=15))
    assert sg["iterations"] == so["iterations"]
    check_solution(w, sg, pg, xg, so, po, xo)


def test_wall_clock_limit_terminates_with_the_state_written_back():
    """max_solver_time_in_seconds (src/params.h:41, src/bundle_adjuster.cpp:18): a tiny limit stops the loop with
    SBA_TERM_MAX_TIME and the last accepted state comes back, like every other termination."""
    w = c2_window()
    args = (w["poses"], w["points"], w["obs_cam"], w["obs_pt"], w["obs_uv"], w["p_l"], w["p_r"])
    sv = Solver(0)
    sv.solve(*args, options=default_options(max_time_s=0.0, max_iterations=2))      # warm-up: module load, pool growth
    pg, xg, sg = sv.solve(*args, options=default_options(max_time_s=2e-3))
    assert sg["termination_name"] == "max_time" and sg["status"] == _capi.SBA_NO_CONVERGENCE
    assert 0 < sg["iterations"] < 50 and sg["final_cost"] < sg["initial_cost"]
    c = oracle.cost(pg, xg, w["obs_cam"], w["obs_pt"], w["obs_uv"], w["p_l"], w["p_r"])
    assert abs(c - sg["final_cost"]) <= 1e-10 * c


def test_numerical_failure_still_writes_back_like_the_reference(small_window):
    """The reference ignores Ceres' termination type and writes back whatever state it has (src/bundle_adjuster.cpp:
    116-132).  A non-finite measurement makes the cost non-finite: SBA_NUMERICAL_FAILURE, the (projected) input state
    is returned instead of an exception."""
    w = small_window
    uv = w["obs_uv"].copy(); uv[5, 0] = np.nan
    pg, xg, sg = Solver(0).solve(w["poses"], w["points"], w["obs_cam"], w["obs_pt"], uv, w["p_l"], w["p_r"], options=PARITY)
    assert sg["status"] == _capi.SBA_NUMERICAL_FAILURE
    assert np.array_equal(pg, w["poses"]) and np.array_equal(xg, w["points"])
