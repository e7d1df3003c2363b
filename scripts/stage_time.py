"""Time one sba_time_stage stage on a configuration: python scripts/stage_time.py C4 5 [reps]"""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
from stereo_orb_slam_b200 import Solver, default_options, pack_window, synthetic
cfg, stage = sys.argv[1], int(sys.argv[2]); reps = int(sys.argv[3]) if len(sys.argv) > 3 else 20
smap = synthetic.make_config(cfg); w = pack_window(smap, 0, smap.num_frames)
sv = Solver(0).set_structure(w["poses"].shape[0], w["points"].shape[0], w["obs_cam"], w["obs_pt"])
sv.set_state(w["poses"], w["points"]); sv.set_measurements(w["obs_uv"], smap.p_l, smap.p_r)
sv.linearize(); sv.schur(1e4); sv.pcg(default_options())
for _ in range(2): sv.time_stage(stage, 3, False, 1e4)
print(cfg, "stage", stage, "%.1f us" % (1e3 * sv.time_stage(stage, reps, False, 1e4)), os.environ.get("SBA_LIB_PATH", "default"))
