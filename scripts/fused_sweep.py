"""Tuning sweep of the fused chunk kernel (occupancy variant x chunk size) on C2: kernel, step and solve times."""
import os, sys, subprocess, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if len(sys.argv) > 1 and sys.argv[1] == "one":
    import numpy as np
    from stereo_orb_slam_b200 import Solver, default_options, pack_window, synthetic
    smap = synthetic.make_config(os.environ.get("SWEEP_CFG", "C2")); w = pack_window(smap, 0, smap.num_frames)
    sv = Solver(0).set_structure(w["poses"].shape[0], w["points"].shape[0], w["obs_cam"], w["obs_pt"])
    sv.set_state(w["poses"], w["points"]); sv.set_measurements(w["obs_uv"], smap.p_l, smap.p_r); sv.linearize()
    for _ in range(3): sv.time_stage(8, 1, True, 1e4)
    out = {"fused_us": 1e3 * sv.time_stage(11, 30, True, 1e4), "step_us": 1e3 * sv.time_stage(8, 30, True, 1e4),
           "lin_us": 1e3 * sv.time_stage(0, 30, True, 1e4), "schur_us": 1e3 * sv.time_stage(2, 30, True, 1e4)}
    s2 = Solver(0); opt = default_options(max_time_s=0.0)
    for _ in range(3):
        p, x, s = s2.solve(w["poses"], w["points"], w["obs_cam"], w["obs_pt"], w["obs_uv"], smap.p_l, smap.p_r, options=opt)
    out.update(solve_ms=1e3 * s["time_solve_s"], its=s["iterations"], dev={k: round(s[k], 2) for k in ("ms_linearize", "ms_schur", "ms_pcg", "ms_update")})
    print(json.dumps(out))
else:
    for env in ({}, {"SBA_OBS_CAP": "120"}):
        e = dict(os.environ); e.update(env)
        r = subprocess.run([sys.executable, __file__, "one"], env=e, capture_output=True, text=True)
        print(env, r.stdout.strip().splitlines()[-1] if r.stdout.strip() else r.stderr[-300:], flush=True)
