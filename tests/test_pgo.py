"""Pose-graph optimisation (reference: src/pose_graph_optimizer.cpp:34-217).  The oracle
(oracle/pgo_oracle.py) restates g2o's EdgeSE3 / VertexSE3 / Huber / Levenberg rules; the GPU path (sba_pgo_optimize)
must follow it trial for trial.  PARITY UNPINNED (g2o is absent from the reference tree and from this image)."""
import numpy as np
import pytest

from oracle import pgo_oracle as P
from stereo_orb_slam_b200 import odometry_graph, optimize_pose_graph, synthetic


def drifting_loop(n, seed, sigma_t=0.02, sigma_r=0.002, outliers=0, loops=1):
    """A closed circuit: ground-truth poses, odometry measurements with drift (so the dead-reckoned start does not
    close), loop-closure edges (id_2 -> id_1 like the reference), optionally a few gross outlier loop edges."""
    rng = np.random.default_rng(seed)
    gt = synthetic.trajectory(n, rng, loop=True)
    Zo = []
    for i in range(n - 1):
        Zi = np.linalg.inv(gt[i]) @ gt[i + 1]
        Zi[:3, :3] = Zi[:3, :3] @ synthetic._rodrigues(rng.normal(0, sigma_r, (1, 3)))[0]
        Zi[:3, 3] += rng.normal(0, sigma_t, 3)
        Zo.append(Zi)
    T0 = np.zeros((n, 4, 4)); T0[0] = gt[0]
    for i in range(n - 1):
        T0[i + 1] = T0[i] @ Zo[i]
    edges = [(i, i + 1) for i in range(n - 1)]
    Z = list(Zo)
    for l in range(loops):            # frame n-1-l re-observes frame l: vertex 0 = id_2 (late), vertex 1 = id_1 (early)
        a, b = n - 1 - l, l
        edges.append((a, b)); Z.append(np.linalg.inv(gt[a]) @ gt[b])
    for o in range(outliers):
        a, b = n // 2 + o, n // 4 + o
        Zb = np.linalg.inv(gt[a]) @ gt[b]; Zb[:3, 3] += rng.normal(0, 30.0, 3)
        edges.append((a, b)); Z.append(Zb)
    # the reference stores poses as float (Eigen::Matrix4f): start from what a float map holds
    return T0.astype(np.float32).astype(np.float64), np.asarray(edges, np.int32), np.asarray(Z, np.float32).astype(np.float64), gt


def test_oracle_jacobians_match_central_differences():
    rng = np.random.default_rng(3)
    for _ in range(5):
        Ri = synthetic._rodrigues(rng.normal(0, 0.6, (1, 3)))[0]; ti = rng.normal(0, 2, 3)
        Rj = synthetic._rodrigues(rng.normal(0, 0.6, (1, 3)))[0]; tj = rng.normal(0, 2, 3)
        Rz = Ri.T @ Rj @ synthetic._rodrigues(rng.normal(0, 0.05, (1, 3)))[0]; tz = Ri.T @ (tj - ti) + rng.normal(0, 0.1, 3)
        e, Ji, Jj = P.edge_jacobians(Ri, ti, Rj, tj, Rz, tz)
        h = 1e-6
        for J, which in ((Ji, 0), (Jj, 1)):
            for k in range(6):
                d = np.zeros(6); d[k] = h
                out = []
                for sgn in (1, -1):
                    dR, dt = P.from_vector_mqt(sgn * d)
                    if which == 0:
                        out.append(P.edge_error(Ri @ dR, ti + Ri @ dt, Rj, tj, Rz, tz)[0])
                    else:
                        out.append(P.edge_error(Ri, ti, Rj @ dR, tj + Rj @ dt, Rz, tz)[0])
                assert np.abs((out[0] - out[1]) / (2 * h) - J[:, k]).max() < 1e-7


def test_oracle_closes_a_drifted_loop():
    T0, edges, Z, gt = drifting_loop(40, seed=1)
    fixed = np.zeros(40, bool); fixed[0] = True
    T1, s = P.optimize(T0, edges, Z, fixed)
    assert s["final_chi2"] < 0.05 * s["initial_chi2"] and s["iterations"] == 10
    gap0 = np.linalg.norm((np.linalg.inv(T0[-1]) @ T0[0])[:3, 3] - (np.linalg.inv(gt[-1]) @ gt[0])[:3, 3])
    gap1 = np.linalg.norm((np.linalg.inv(T1[-1]) @ T1[0])[:3, 3] - (np.linalg.inv(gt[-1]) @ gt[0])[:3, 3])
    assert gap1 < 0.2 * gap0
    e, Zo = odometry_graph(gt)
    assert e.shape == (39, 2) and np.allclose(gt[3] @ Zo[3], gt[4])


@pytest.mark.gpu
@pytest.mark.parametrize("n,outliers,loops", [(12, 0, 1), (120, 0, 2), (400, 3, 4)])
def test_gpu_pose_graph_follows_the_oracle(n, outliers, loops):
    T0, edges, Z, gt = drifting_loop(n, seed=10 + n, outliers=outliers, loops=loops)
    fixed = np.zeros(n, np.uint8); fixed[0] = 1
    Tg, sg = optimize_pose_graph(T0, edges, Z)
    To, so = P.optimize(T0, edges, Z, fixed.astype(bool))
    # once converged, the gain of a step is rounding noise and a trial may be accepted on one side and retried on the
    # other: iterations agree, trials within a few
    assert sg["iterations"] == so["iterations"] and abs(sg["trials"] - so["trials"]) <= 3
    assert abs(sg["initial_chi2"] - so["initial_chi2"]) <= 1e-12 * so["initial_chi2"]
    assert abs(sg["final_chi2"] - so["final_chi2"]) <= 1e-7 * max(so["final_chi2"], 1e-12)
    assert np.abs(Tg[:, :3, :3] - To[:, :3, :3]).max() < 1e-8 and np.abs(Tg[:, :3, 3] - To[:, :3, 3]).max() < 1e-7
    assert np.array_equal(Tg[0], T0[0])                                    # vertex 0 fixed (:123-126)
    if outliers:
        assert sg["final_chi2"] > 1.0                                      # the Huber kernel kept the gross edges from being fitted


@pytest.mark.gpu
def test_gpu_pose_graph_partial_optimisation_and_custom_information():
    """Partial PGO (src/pose_graph_optimizer.cpp:45-54): only a range of frames is free."""
    n = 60
    T0, edges, Z, gt = drifting_loop(n, seed=77, loops=1)
    fixed = np.zeros(n, np.uint8); fixed[:20] = 1
    info = np.array([0.5, 0.5, 0.5, 2.0, 2.0, 2.0])
    Tg, sg = optimize_pose_graph(T0, edges, Z, fixed=fixed, information=info, huber_delta=0.7, iterations=6)
    To, so = P.optimize(T0, edges, Z, fixed.astype(bool), iterations=6, info=np.diag(info), delta=0.7)
    assert np.array_equal(Tg[:20], T0[:20]) and sg["iterations"] == so["iterations"] and abs(sg["trials"] - so["trials"]) <= 3
    assert abs(sg["final_chi2"] - so["final_chi2"]) <= 1e-7 * so["final_chi2"] and sg["final_chi2"] < sg["initial_chi2"]
    assert np.abs(Tg - To).max() < 1e-7
