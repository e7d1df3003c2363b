import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from stereo_orb_slam_b200 import Solver, pack_window, synthetic
smap = synthetic.make_config("C2"); w = pack_window(smap, 0, smap.num_frames)
sv = Solver(0).set_structure(w["poses"].shape[0], w["points"].shape[0], w["obs_cam"], w["obs_pt"])
sv.set_state(w["poses"], w["points"]); sv.set_measurements(w["obs_uv"], smap.p_l, smap.p_r)
sv.linearize(); sv.schur(1e4)
for flush in (True, False):
    for name, m in [("full", 0), ("schur -main", 1), ("schur -rhs", 4), ("schur -main-rhs", 5), ("schur stage only", 8), ("lin -slot", 1 << 8), ("lin -point", 4 << 8), ("lin -slot-point", 5 << 8)]:
        os.environ["SBA_DEBUG_MASK"] = str(m)
        print("flush=%d %-24s lin %.1f us  schur %.1f us" % (flush, name, 1e3 * sv.time_stage(0, 20, flush, 1e4), 1e3 * sv.time_stage(2, 20, flush, 1e4)))
