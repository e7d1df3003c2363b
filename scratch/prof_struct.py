import sys, time, numpy as np
sys.path.insert(0, '.')
from stereo_orb_slam_b200 import Solver, pack_window, synthetic
name = sys.argv[1] if len(sys.argv) > 1 else "C2"
scale = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
t=time.time(); smap = synthetic.make_config(name, scale=scale); print("gen", time.time()-t)
t=time.time(); w = pack_window(smap, 0, smap.num_frames); print("pack", time.time()-t, w["obs_cam"].size)
s = Solver(-1)
for i in range(3):
    t=time.time(); s.set_structure(w["poses"].shape[0], w["points"].shape[0], w["obs_cam"], w["obs_pt"]); print("structure", time.time()-t)
print("blocks", s.schur_pattern()[0].size)
