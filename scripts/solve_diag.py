import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from stereo_orb_slam_b200 import Solver, default_options, pack_window, synthetic
for name in sys.argv[1:] or ["C1", "C2"]:
    smap = synthetic.make_config(name)
    w = pack_window(smap, 0, smap.num_frames)
    sv = Solver(0)
    opt = default_options(max_time_s=0.0)
    for rep in range(2):
        t0 = time.perf_counter()
        p, x, s = sv.solve(w["poses"], w["points"], w["obs_cam"], w["obs_pt"], w["obs_uv"], smap.p_l, smap.p_r, options=opt)
        dt = time.perf_counter() - t0
    print(name, "obs", w["obs_cam"].size, "it", s["iterations"], "succ", s["successful_steps"], "lin", s["linearizations"],
          "pcg_its", s["pcg_iterations"], "term", s["termination_name"], "cost %.9e -> %.9e" % (s["initial_cost"], s["final_cost"]))
    print("   wall %.2f ms: structure %.2f h2d %.2f solve %.2f d2h %.2f | dev ms lin %.2f schur %.2f pcg %.2f update %.2f | launches %d" % (
        1e3 * dt, 1e3 * s["time_structure_s"], 1e3 * s["time_h2d_s"], 1e3 * s["time_solve_s"], 1e3 * s["time_d2h_s"],
        s["ms_linearize"], s["ms_schur"], s["ms_pcg"], s["ms_update"], sv.launch_count))
