import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from stereo_orb_slam_b200 import Solver, pack_window, synthetic
name = sys.argv[1] if len(sys.argv) > 1 else "C2"
smap = synthetic.make_config(name); w = pack_window(smap, 0, smap.num_frames)
sv = Solver(0).set_structure(w["poses"].shape[0], w["points"].shape[0], w["obs_cam"], w["obs_pt"])
sv.set_state(w["poses"], w["points"]); sv.set_measurements(w["obs_uv"], smap.p_l, smap.p_r)
sv.linearize(); sv.schur(1e4); sv.pcg()
names = {0: "k_lin_chunk", 9: "k_assemble_cams", 10: "k_lin_finish", 3: "linearize", 2: "k_schur_chunk", 1: "k_assemble_schur", 4: "schur", 8: "lin+schur", 5: "k_backsub", 6: "k_trial_points", 7: "pcg iteration (multi-launch)"}
for flush in (True, False):
    print("flush", flush, " ".join("%s %.1f" % (names[k], 1e3 * sv.time_stage(k, 20, flush, 1e4)) for k in (0, 9, 10, 3, 2, 1, 4, 8, 5, 6)))
