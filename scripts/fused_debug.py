"""Where do the fused and the two-kernel LM trajectories part?  (debug aid for tests/test_gpu_parity.py::test_fused_...)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from stereo_orb_slam_b200 import Solver, default_options, pack_window, synthetic
smap = synthetic.make_map(num_frames=8, num_points=600, track_len=5, seed=7)
w = pack_window(smap, 0, 8)
args = (w["poses"], w["points"], w["obs_cam"], w["obs_pt"], w["obs_uv"], smap.p_l, smap.p_r)
res = {}
for name, env in (("fused", {}), ("unfused", {"SBA_NO_FUSED": "1"}), ("fused_nofork", {"SBA_NO_FORK": "1"}), ("unfused_nofork", {"SBA_NO_FUSED": "1", "SBA_NO_FORK": "1"})):
    for k in ("SBA_NO_FUSED", "SBA_NO_FORK"):
        os.environ.pop(k, None)
    os.environ.update(env)
    costs = []
    for it in range(1, 12):
        p, x, s = Solver(0).solve(*args, options=default_options(max_time_s=0.0, max_iterations=it))
        costs.append(s["final_cost"])
    res[name] = (costs, p, x)
    print(name, ["%.17g" % c for c in costs[:6]])
for a, b in (("fused", "unfused"), ("fused", "fused_nofork"), ("unfused", "unfused_nofork")):
    d = [abs(u - v) / v for u, v in zip(res[a][0], res[b][0])]
    print(a, "vs", b, "first differing iteration:", next((i + 1 for i, e in enumerate(d) if e != 0), None), "max rel", max(d))
