"""Timing of the stereo RANSAC alignment (N4) through the C-ABI with host buffers: one frame pair (the visual
odometer's per-frame call) and a batch of independent problems; the numpy oracle is timed beside it on the host.
Usage: python scripts/ransac_bench.py [--n 500] [--batch 1024] > profiles/rNN_ransac.json"""
import argparse, json, os, sys, time

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
from stereo_orb_slam_b200 import synthetic                      # noqa: E402
from stereo_orb_slam_b200.ransac import GlibcRand, StereoRansac  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=500)
    ap.add_argument("--batch", type=int, default=1024)
    ap.add_argument("--reps", type=int, default=50)
    a = ap.parse_args()
    pairs, T_true, bad, P_l, P_r = synthetic.make_stereo_pairs(a.n, 1)
    solver = StereoRansac(P_l, P_r)
    S = GlibcRand(1).draws(100, a.n)
    for _ in range(5):
        solver.Solve(pairs, samples=S)
    L0 = solver._sv.launch_count
    ts = []
    for _ in range(a.reps):
        t = time.perf_counter(); T, labels, losses = solver.Solve(pairs, samples=S); ts.append(time.perf_counter() - t)
    launches = (solver._sv.launch_count - L0) / a.reps
    out = dict(single=dict(n=a.n, hypotheses=100, ms_median=1e3 * float(np.median(ts)), ms_min=1e3 * float(np.min(ts)),
                           inlier_tests_per_s=100 * a.n / float(np.median(ts)), kernel_launches_per_call=launches,
                           mask_accuracy=float((labels == ~bad).mean())))
    from oracle import ransac_oracle as RO                      # checker / CPU baseline leg only
    t = time.perf_counter(); To, lo, so, info = RO.solve(pairs[:, 0], pairs[:, 1], P_l, P_r, S); dt = time.perf_counter() - t
    out["cpu_oracle"] = dict(kind="port", cores=1, ms=1e3 * dt, inlier_tests_per_s=100 * a.n / dt,
                             note="numpy restatement, triangulation hoisted out of the hypothesis loop (the reference re-triangulates per test)",
                             same_best_iteration=bool(info["best_iteration"] == solver.last["best_iteration"]),
                             max_abs_T_diff=float(np.abs(To - T).max()), label_mismatches=int((lo != labels).sum()))
    rng = np.random.default_rng(0)
    firsts = [pairs[:, 0]] * a.batch; seconds = [pairs[:, 1]] * a.batch
    samples = [rng.integers(0, a.n, (100, 3)).astype(np.int32) for _ in range(a.batch)]
    solver.solve_batch(firsts, seconds, samples)
    ts = []
    for _ in range(5):
        t = time.perf_counter(); solver.solve_batch(firsts, seconds, samples); ts.append(time.perf_counter() - t)
    out["batch"] = dict(problems=a.batch, n=a.n, ms_median=1e3 * float(np.median(ts)),
                        problems_per_s=a.batch / float(np.median(ts)), inlier_tests_per_s=100.0 * a.n * a.batch / float(np.median(ts)),
                        note="wall clock of the Python call: includes concatenating the host arrays")
    print(json.dumps(out))


if __name__ == "__main__":
    main()
