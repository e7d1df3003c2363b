import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from stereo_orb_slam_b200 import Solver, default_options, pack_window, synthetic
scale = float(sys.argv[1]) if len(sys.argv) > 1 else 1.0
smap = synthetic.make_config("C4", scale=scale)
b = pack_window(smap, 0, smap.num_frames)
sv = Solver(0)
sv.set_structure(b["poses"].shape[0], b["points"].shape[0], b["obs_cam"], b["obs_pt"])
sv.set_state(b["poses"], b["points"]); sv.set_measurements(b["obs_uv"], smap.p_l, smap.p_r)
def lin():
    sv.linearize(); g = sv.get_linearization(False)
    return {k: np.array(v, copy=True) for k, v in g.items() if k in ("U", "V", "gc", "gp", "cost", "W")}
a = lin(); c = lin()
for k in a: print("linearize", k, "same bits:", np.array_equal(a[k], c[k]), flush=True)
def sch():
    sv.schur(1e4); bl, rhs, _ = sv.schur_values(); return bl.copy(), rhs.copy()
s1 = sch(); s2 = sch()
print("schur blocks same:", np.array_equal(s1[0], s2[0]), "rhs same:", np.array_equal(s1[1], s2[1]), "| max block diff", np.abs(s1[0]-s2[0]).max(), flush=True)
def pcg():
    it = sv.pcg(default_options(max_pcg_iterations=500)); return it, sv.schur_values()[2].copy()
p1 = pcg(); p2 = pcg()
print("pcg iterations", p1[0], p2[0], "x same:", np.array_equal(p1[1], p2[1]), "max diff", np.abs(p1[1]-p2[1]).max(), "max |x|", np.abs(p1[1]).max(), flush=True)
