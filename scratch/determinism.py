import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from stereo_orb_slam_b200 import Solver, default_options, pack_window, synthetic
name = sys.argv[1] if len(sys.argv) > 1 else "C4"; scale = float(sys.argv[2]) if len(sys.argv) > 2 else 0.25
smap = synthetic.make_config(name, scale=scale) if name == "C4" else synthetic.make_config(name)
b = pack_window(smap, 0, smap.num_frames)
opt = default_options(max_time_s=0.0, max_iterations=10)
res = []
for rep in range(3):
    sv = Solver(0)
    sv.set_structure(b["poses"].shape[0], b["points"].shape[0], b["obs_cam"], b["obs_pt"])
    sv.set_state(b["poses"], b["points"]); sv.set_measurements(b["obs_uv"], smap.p_l, smap.p_r)
    if rep == 2:
        sv.linearize()
        for _ in range(5): sv.time_stage(8, 1, True, 1e4)
        sv.set_state(b["poses"], b["points"])
    s = sv.run(opt)
    poses, pts = sv.get_state()
    res.append((s["final_cost"], s["iterations"], s["pcg_iterations"], poses.copy(), pts.copy()))
    print(name, "rep", rep, "cost %.12e it %d pcg %d" % (s["final_cost"], s["iterations"], s["pcg_iterations"]), flush=True)
    sv.close()
print("rep0 == rep1 bits:", np.array_equal(res[0][3], res[1][3]) and np.array_equal(res[0][4], res[1][4]),
      "| rep0 == rep2 (after timing stages):", np.array_equal(res[0][3], res[2][3]) and np.array_equal(res[0][4], res[2][4]))
