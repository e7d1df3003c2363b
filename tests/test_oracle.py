"""The oracle against the golden vectors (tests/golden, mpmath 50 digits) and against itself.

Parity is UNPINNED by the reference (it has no tests; Ceres is absent) — these KATs pin the
residual / Jacobian / Huber arithmetic of reference: src/reprojection_error.h:12-41,
src/bundle_adjuster.cpp:100 independently of both oracle/ and csrc/.
"""
import json
import os

import numpy as np
import pytest

from oracle import oracle
from conftest import make_window

HERE = os.path.dirname(os.path.abspath(__file__))
KATS = json.load(open(os.path.join(HERE, "golden", "reprojection_kats.json")))


@pytest.mark.parametrize("case", KATS["cases"], ids=[c["name"] for c in KATS["cases"]])
@pytest.mark.parametrize("mode", [0, 1], ids=["jet", "analytic"])
def test_residual_jacobian_kat(case, mode):
    r, Jc, Jp = oracle.residual(case["cam"], case["point"], case["uv"], case["p_l"], case["p_r"], mode)
    assert np.abs(r - np.array(case["r"])).max() < 1e-11          # pixels
    # Jet autodiff differentiates Ceres's first-order small-angle branch: exact to O(theta) only
    tol = 1e-8 if mode == 0 else 1e-12
    Jc_ref, Jp_ref = np.array(case["Jc"]), np.array(case["Jp"])
    assert np.abs(Jc - Jc_ref).max() <= tol * np.abs(Jc_ref).max()
    assert np.abs(Jp - Jp_ref).max() <= 1e-12 * np.abs(Jp_ref).max()


@pytest.mark.parametrize("h", KATS["huber"], ids=[str(h["s"]) for h in KATS["huber"]])
def test_huber_kat(h):
    rho, sq = oracle.huber(1.0, h["s"])
    assert abs(rho - h["rho"]) <= 1e-15 * max(1.0, abs(h["rho"]))
    assert abs(sq - h["sqrt_rho1"]) <= 1e-15


def test_device_formulas_match_oracle_blocks(emu, small_window):
    """csrc/sba_math.cuh (compiled for the host) reproduces the oracle's per-observation blocks."""
    import ctypes as C
    w = small_window
    lin = oracle.linearize(w["poses"], w["points"], w["obs_cam"], w["obs_pt"], w["obs_uv"], w["p_l"], w["p_r"])
    d = C.c_double
    P = lambda a, t=d: a.ctypes.data_as(C.POINTER(t))
    Nc, Np = w["poses"].shape[0], w["points"].shape[0]
    U = np.zeros((Nc, 6, 6)); V = np.zeros((Np, 3, 3)); gc = np.zeros((Nc, 6)); gp = np.zeros((Np, 3)); cost = 0.0
    iu = np.triu_indices(6)
    for k in range(w["obs_cam"].size):
        c, p = int(w["obs_cam"][k]), int(w["obs_pt"][k])
        r = np.zeros(4); U21 = np.zeros(21); g6 = np.zeros(6); V6 = np.zeros(6); g3 = np.zeros(3); W = np.zeros(18); co = d(0)
        pose = np.ascontiguousarray(w["poses"][c]); X = np.ascontiguousarray(w["points"][p]); uv = np.ascontiguousarray(w["obs_uv"][k])
        emu.emu_observation(P(pose), P(X), P(uv, C.c_float), P(w["p_l"]), P(w["p_r"]), 4, d(1.0), P(r), P(U21), P(g6), P(V6),
                            P(g3), P(W), C.byref(co))
        assert np.abs(r - lin["r"][k]).max() < 1e-9
        assert np.abs(W.reshape(6, 3) - lin["W"][k]).max() <= 1e-9 * max(1.0, np.abs(lin["W"][k]).max())
        Uf = np.zeros((6, 6)); Uf[iu] = U21; Uf = Uf + np.triu(Uf, 1).T
        Vf = np.array([[V6[0], V6[1], V6[2]], [V6[1], V6[3], V6[4]], [V6[2], V6[4], V6[5]]])
        U[c] += Uf; V[p] += Vf; gc[c] += g6; gp[p] += g3; cost += co.value
    for a, b in ((U, lin["U"]), (V, lin["V"]), (gc, lin["gc"]), (gp, lin["gp"])):
        assert np.abs(a - b).max() <= 1e-9 * np.abs(b).max()
    assert abs(cost - lin["cost"]) <= 1e-12 * lin["cost"]


def test_oracle_noise_free_recovers_ground_truth():
    """Known-answer case: exact measurements => the minimum is the ground truth, cost -> ~0
    (limited by the f32 storage of the pixels)."""
    w = make_window(6, 300, 4, seed=11, noise_free=True)
    smap = w["map"]
    po, xo, s = oracle.solve(w["poses"], w["points"], w["obs_cam"], w["obs_pt"], w["obs_uv"], w["p_l"], w["p_r"],
                             options=oracle.default_options(max_time_s=0.0, num_threads=2))
    assert s["termination"] == 0
    assert s["final_cost"] < 1e-4 * s["initial_cost"] and s["final_cost"] / w["obs_cam"].size < 1e-6
    # gauge: camera 0 is held at its (exact) initial pose, so the solution is the ground truth itself
    gt_pts = smap.gt_points[w["point_ids"]]
    assert np.abs(xo - gt_pts).max() < 2e-2   # metres; far points are weakly constrained by 1e-4 px f32 rounding
    assert np.median(np.abs(xo - gt_pts)) < 1e-3


def test_oracle_minimum_is_solver_independent(small_window):
    """autodiff and analytic Jacobians, with and without the line search, reach the same minimum."""
    w = small_window
    res = []
    for kw in (dict(), dict(analytic_jacobian=1), dict(line_search=0)):
        po, xo, s = oracle.solve(w["poses"], w["points"], w["obs_cam"], w["obs_pt"], w["obs_uv"], w["p_l"], w["p_r"],
                                 options=oracle.default_options(max_time_s=0.0, num_threads=2, **kw))
        res.append((po, xo, s))
    for po, xo, s in res[1:]:
        assert abs(s["final_cost"] - res[0][2]["final_cost"]) <= 1e-7 * res[0][2]["final_cost"]
        assert np.abs(po - res[0][0]).max() < 1e-5 and np.abs(xo - res[0][1]).max() < 1e-4


def test_oracle_vs_numpy_gradient(small_window):
    """The oracle's gradient (J^T r with the Huber corrector) equals the numerical gradient of
    E = 1/2 sum rho(||r||^2) evaluated by an independent numpy restatement of the objective."""
    w = small_window
    pl, pr = w["p_l"].reshape(3, 4), w["p_r"].reshape(3, 4)

    def objective(poses, pts):
        th = np.maximum(np.linalg.norm(poses[:, :3], axis=1), 1e-150)[:, None, None]
        K = np.zeros((poses.shape[0], 3, 3))
        K[:, 0, 1] = -poses[:, 2]; K[:, 0, 2] = poses[:, 1]; K[:, 1, 0] = poses[:, 2]
        K[:, 1, 2] = -poses[:, 0]; K[:, 2, 0] = -poses[:, 1]; K[:, 2, 1] = poses[:, 0]
        R = np.eye(3) + np.sin(th) / th * K + (1 - np.cos(th)) / th ** 2 * (K @ K)
        c, p = w["obs_cam"], w["obs_pt"]
        q = np.einsum("kij,kj->ki", R[c], pts[p]) + poses[c, 3:]
        qh = np.concatenate([q, np.ones((q.shape[0], 1))], axis=1)
        a, b = qh @ pl.T, qh @ pr.T
        pred = np.stack([a[:, 0] / a[:, 2], a[:, 1] / a[:, 2], b[:, 0] / b[:, 2], b[:, 1] / b[:, 2]], axis=1)
        s = ((pred - w["obs_uv"].astype(np.float64)) ** 2).sum(axis=1)
        return 0.5 * np.where(s <= 1.0, s, 2 * np.sqrt(s) - 1).sum()

    lin = oracle.linearize(w["poses"], w["points"], w["obs_cam"], w["obs_pt"], w["obs_uv"], w["p_l"], w["p_r"])
    assert abs(objective(w["poses"], w["points"]) - lin["cost"]) <= 1e-10 * lin["cost"]
    rng = np.random.default_rng(3)
    for _ in range(6):
        c, i = int(rng.integers(1, w["poses"].shape[0])), int(rng.integers(0, 6))
        h = 1e-6
        pp = w["poses"].copy(); pp[c, i] += h; pm = w["poses"].copy(); pm[c, i] -= h
        num = (objective(pp, w["points"]) - objective(pm, w["points"])) / (2 * h)
        assert abs(num - lin["gc"][c, i]) <= 1e-5 * max(1.0, abs(num))
    for _ in range(6):
        p, i = int(rng.integers(0, w["points"].shape[0])), int(rng.integers(0, 3))
        h = 1e-6
        xp = w["points"].copy(); xp[p, i] += h; xm = w["points"].copy(); xm[p, i] -= h
        num = (objective(w["poses"], xp) - objective(w["poses"], xm)) / (2 * h)
        assert abs(num - lin["gp"][p, i]) <= 1e-5 * max(1.0, abs(num))


def test_line_search_step_selection_three_ways(emu):
    """Ceres's MinimizeInterpolatingPolynomial restated three times — oracle (C++, Laguerre roots), the device
    header csrc/sba_linesearch.cuh (Durand-Kerner roots, compiled for the host) and numpy (polyfit-free
    Vandermonde solve + np.roots) — must pick the same step, incl. the real parts of complex roots."""
    import ctypes as C
    rng = np.random.default_rng(8)
    d = C.c_double
    P = lambda a, t=d: a.ctypes.data_as(C.POINTER(t))
    for trial in range(200):
        n = 2 if trial % 2 == 0 else 3
        x = np.zeros(n); f = np.zeros(n); g = np.zeros(n)
        f[0] = rng.uniform(1, 100); g[0] = -rng.uniform(0.1, 50)
        x[1] = rng.uniform(1e-3, 1.0); f[1] = f[0] + rng.normal(0, 5); g[1] = rng.normal(0, 50)
        if n == 3:
            x[2] = x[1] / rng.uniform(0.05, 0.6); f[2] = f[0] + abs(rng.normal(0, 20)); g[2] = rng.normal(0, 80)
        hg = np.ones(n, np.int32)
        lo, hi = 1e-3 * x[1], 0.6 * x[1]
        a = oracle.ls_minimize(x, f, g, hg, lo, hi)
        b = emu.emu_ls_minimize(n, P(x), P(f), P(g), P(hg, C.c_int), d(lo), d(hi))
        # numpy restatement
        deg = 2 * n - 1
        rows, rhs = [], []
        for i in range(n):
            rows.append([x[i] ** (deg - j) for j in range(deg + 1)]); rhs.append(f[i])
            rows.append([(deg - j) * x[i] ** (deg - j - 1) if j < deg else 0.0 for j in range(deg + 1)]); rhs.append(g[i])
        c = np.linalg.solve(np.array(rows), np.array(rhs))
        cand = [0.5 * (lo + hi), lo, hi] + [float(np.real(r)) for r in np.roots(np.polyder(c))]
        cand = [v for v in cand if lo <= v <= hi]
        vals = [np.polyval(c, v) for v in cand]
        ref = cand[int(np.argmin(vals))]
        pa, pb, pr = np.polyval(c, a), np.polyval(c, b), min(vals)
        scale = max(1.0, abs(pr))
        assert abs(pa - pr) <= 1e-9 * scale and abs(pb - pr) <= 1e-9 * scale, (trial, a, b, ref)
        assert abs(a - b) <= 1e-7 * hi, (trial, a, b, ref)


def test_polynomial_root_real_parts_match_numpy(emu):
    """sba_linesearch.cuh's root finder (closed form up to degree 2, Laguerre + deflation above) against
    numpy.roots, including the badly scaled case the line search produces (tiny leading coefficient)."""
    import ctypes as C
    rng = np.random.default_rng(11)
    cases = [rng.normal(size=d + 1) for d in (1, 2, 3, 4) for _ in range(20)]
    cases += [np.array([1e-14, 2.0, -3.0, 0.5, 1.0]), np.array([1e-9, 1e-3, 1.0, -2.0]), np.array([1.0, -4.0, 6.0, -4.0, 1.0 - 1e-8]),
              np.array([1.0, 0.0, 2.0, 0.0, 1.0 + 1e-6])]
    for c in cases:
        c = np.ascontiguousarray(c, np.float64)
        out = np.zeros(4)
        n = emu.emu_ls_roots(c.ctypes.data_as(C.POINTER(C.c_double)), C.c_int(c.size - 1), out.ctypes.data_as(C.POINTER(C.c_double)))
        assert n == c.size - 1
        ref = np.sort(np.roots(c).real)
        got = np.sort(out[:n])
        scale = np.maximum(1.0, np.abs(np.roots(c)).max())
        assert np.abs(got - ref).max() <= 1e-6 * scale, (c, got, ref)


def test_envelope_cholesky_path_equals_dense_path(monkeypatch):
    """Global problems go through a row-envelope Cholesky instead of the dense LLT (oracle/sba_oracle.cpp `Envelope`):
    same reduced system, same solution — forced both ways on an open chain with loop-closure-free band structure and
    on a loop, the trajectories must agree to rounding."""
    from stereo_orb_slam_b200 import pack_window, synthetic
    for smap in (synthetic.make_config("C5", scale=0.003), synthetic.make_config("C4", scale=0.015)):
        w = pack_window(smap, 0, smap.num_frames)
        out = {}
        for env in ("0", "1"):
            monkeypatch.setenv("ORACLE_ENVELOPE", env)
            out[env] = oracle.solve(w["poses"], w["points"], w["obs_cam"], w["obs_pt"], w["obs_uv"], smap.p_l, smap.p_r,
                                    options=oracle.default_options(max_time_s=0.0, num_threads=0, max_iterations=12))
        assert out["0"][2]["iterations"] == out["1"][2]["iterations"] == 12
        assert abs(out["0"][2]["final_cost"] - out["1"][2]["final_cost"]) <= 1e-11 * out["0"][2]["final_cost"]
        assert np.abs(out["0"][0] - out["1"][0]).max() < 1e-9 and np.abs(out["0"][1] - out["1"][1]).max() < 1e-8
