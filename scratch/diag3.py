import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
from stereo_orb_slam_b200 import Solver, default_options, pack_window, synthetic
def run(name, kw, radius):
    smap = synthetic.make_config(name) if name else synthetic.make_map(**kw)
    w = pack_window(smap, 0, smap.num_frames)
    sv = Solver(0).set_structure(w["poses"].shape[0], w["points"].shape[0], w["obs_cam"], w["obs_pt"])
    sv.set_state(w["poses"], w["points"]); sv.set_measurements(w["obs_uv"], smap.p_l, smap.p_r)
    sv.linearize(); sv.schur(radius)
    it = sv.pcg()
    S, rhs, x = sv.schur_dense()
    xs = np.linalg.solve(S, rhs)
    print(name or kw, "radius", radius, "pcg its", it, "rel err", np.abs(x - xs).max() / np.abs(xs).max(), "cond", np.linalg.cond(S))
run(None, dict(num_frames=8, num_points=600, track_len=5, seed=7), 1e4)
run("C2", None, 1e4)
run("C2", None, 1e8)
run("C2", None, 1e12)
