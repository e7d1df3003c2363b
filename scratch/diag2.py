import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from stereo_orb_slam_b200 import Solver, default_options, pack_window, synthetic
from oracle import oracle
smap = synthetic.make_config("C2")
w = pack_window(smap, 0, smap.num_frames)
sv = Solver(0)
opt = default_options(max_time_s=0.0, verbose=1, max_iterations=int(sys.argv[1]) if len(sys.argv) > 1 else 60, pcg_tolerance=float(os.environ.get('PCG_TOL', '1e-10')), max_pcg_iterations=2000)
p, x, s = sv.solve(w["poses"], w["points"], w["obs_cam"], w["obs_pt"], w["obs_uv"], smap.p_l, smap.p_r, options=opt)
print("GPU", s["iterations"], s["termination_name"], "%.12e" % s["final_cost"], "pcg", s["pcg_iterations"])
if os.environ.get("WITH_ORACLE"):
    po, xo, so = oracle.solve(w["poses"], w["points"], w["obs_cam"], w["obs_pt"], w["obs_uv"], smap.p_l, smap.p_r,
                              options=oracle.default_options(max_time_s=0.0, num_threads=0, max_iterations=opt.max_iterations, verbose=1))
    print("ORACLE", so["iterations"], so.get("termination_name", so.get("termination")), "%.12e" % so["final_cost"])
    print("max dpose", np.abs(p - po).max(), "max dX", np.abs(x - xo).max())
