import sys, numpy as np
sys.path.insert(0, '.')
from stereo_orb_slam_b200 import Solver, default_options, pack_window, synthetic
from oracle import oracle
smap = synthetic.make_map(num_frames=7, num_points=300, track_len=4, seed=17)
pid = smap.obs_point_id.copy(); off = smap.frame_off
pid[off[3] + 1] = pid[off[3]]; pid[off[4] + 5] = pid[off[4] + 4]
smap.obs_point_id = pid
w = pack_window(smap, 0, 7)
poses, points, s = Solver(0).solve(w["poses"], w["points"], w["obs_cam"], w["obs_pt"], w["obs_uv"], smap.p_l, smap.p_r,
                                   options=default_options(max_time_s=0.0, verbose=1))
print(s)
po, xo, so = oracle.solve(w["poses"], w["points"], w["obs_cam"], w["obs_pt"], w["obs_uv"], smap.p_l, smap.p_r,
                          options=oracle.default_options(max_time_s=0.0, num_threads=0, verbose=1))
print(so)
