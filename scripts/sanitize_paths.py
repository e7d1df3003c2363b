"""Every solver path on a small problem, TWICE, results compared bit for bit.

Written as the workload for compute-sanitizer (memcheck / racecheck):
  compute-sanitizer --tool memcheck  python scripts/sanitize_paths.py
On this pool compute-sanitizer is closed (profiles/r02_sanitizer_closed.log: "runs under it have left GPUs needing a
reset"), so the script is its own check: every reduction in the library has a fixed order, hence any data race shows
up as a run-to-run difference in the last bits (that is how round 1's k_pcg_coop race was found); each path must
reproduce itself exactly.

Paths: chunk kernels fused and unfused, k_chol_band / k_chol_window / k_chol_cluster, k_spchol (panel and column
paths), the four PCG paths, k_step_loop and the multi-launch step, long tracks, batched windows, the resident map."""
import os, sys, subprocess
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))

CASES = [
    ("band cholesky, fused", {}, 0),
    ("band cholesky, two-kernel linearisation", {"SBA_NO_FUSED": "1"}, 0),
    ("window cholesky", {"SBA_NO_BAND": "1", "SBA_NO_CLUSTER": "1"}, 0),
    ("cluster cholesky", {"SBA_FORCE_CLUSTER": "1"}, 0),
    ("sparse cholesky (forced)", {"SBA_FORCE_SPARSE_CHOL": "1"}, 0),
    ("pcg window", {"SBA_NO_CLUSTER": "1"}, 1),
    ("pcg cluster", {"SBA_FORCE_CLUSTER": "1"}, 1),
    ("pcg multilaunch", {"SBA_FORCE_MULTILAUNCH": "1"}, 1),
    ("pcg cooperative", {"SBA_FORCE_COOP_PCG": "1"}, 1),
    ("multi-launch step (no cooperative launch)", {"SBA_NO_COOP": "1"}, 0),
]


def one(kind, solver):
    import numpy as np
    from stereo_orb_slam_b200 import Solver, default_options, pack_window, replay, synthetic
    opt = default_options(max_time_s=0.0, max_iterations=6, linear_solver=solver)
    if kind == "small":
        smap = synthetic.make_map(num_frames=8, num_points=600, track_len=5, seed=7)
    elif kind == "loop":
        smap = synthetic.make_config("C4", scale=0.03)
    elif kind == "long":
        from conftest import long_track_map
        smap = long_track_map(num_frames=258, num_long=2, num_short=80)
    elif kind == "map":
        smap = synthetic.make_map(num_frames=12, num_points=500, track_len=5, seed=11)
        outs = []
        for _ in range(2):
            ba, calls = replay(smap, interval=5, options=opt)
            outs.append((ba.poses().copy(), ba.points().copy()))
        same = np.array_equal(outs[0][0], outs[1][0]) and np.array_equal(outs[0][1], outs[1][1])
        print("   map replay:", len(calls), "calls, last cost %.6e, two replays %s" % (calls[-1][2]["final_cost"], "bit-identical" if same else "DIFFER"))
        if not same:
            raise SystemExit("run-to-run difference in the map replay")
        return
    w = pack_window(smap, 0, smap.num_frames)
    runs = [Solver(0).solve(w["poses"], w["points"], w["obs_cam"], w["obs_pt"], w["obs_uv"], smap.p_l, smap.p_r, options=opt) for _ in range(3)]
    p, x, s = runs[0]
    same = all(np.array_equal(r[0], p) and np.array_equal(r[1], x) and r[2]["final_cost"] == s["final_cost"] for r in runs[1:])
    print("   %s: cost %.6e -> %.6e, %d iterations, three runs %s" % (kind, s["initial_cost"], s["final_cost"], s["iterations"],
                                                                       "bit-identical" if same else "DIFFER"))
    if not same:
        raise SystemExit("run-to-run difference on path: " + kind)


if __name__ == "__main__":
    if len(sys.argv) > 1:
        one(sys.argv[1], int(sys.argv[2]))
        sys.exit(0)
    # every case in THIS process (the sanitizer follows it): the environment switches are read at set_structure / launch
    for name, env, solver in CASES:
        for k in ("SBA_NO_FUSED", "SBA_NO_BAND", "SBA_NO_CLUSTER", "SBA_FORCE_CLUSTER", "SBA_FORCE_SPARSE_CHOL", "SBA_FORCE_MULTILAUNCH",
                  "SBA_FORCE_COOP_PCG", "SBA_NO_COOP"):
            os.environ.pop(k, None)
        os.environ.update(env)
        print(name, flush=True)
        one("small", solver)
    for k in ("SBA_NO_FUSED", "SBA_NO_BAND", "SBA_NO_CLUSTER", "SBA_FORCE_CLUSTER", "SBA_FORCE_SPARSE_CHOL", "SBA_FORCE_MULTILAUNCH",
              "SBA_FORCE_COOP_PCG", "SBA_NO_COOP"):
        os.environ.pop(k, None)
    print("sparse cholesky on a loop (panel path)", flush=True); one("loop", 0)
    print("long tracks", flush=True); one("long", 0)
    print("resident map", flush=True); one("map", 0)
    print("done")
