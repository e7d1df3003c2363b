"""Host structure-build phases (SBA_TIMING=1) on a global problem, unsharded and as rank 0 of an 8-way point shard.
Usage: SBA_TIMING=1 python scripts/struct_prof.py [C4] [threads]"""
import os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
import numpy as np
cfg = sys.argv[1] if len(sys.argv) > 1 else "C4"
if len(sys.argv) > 2:
    os.environ["SBA_HOST_THREADS"] = sys.argv[2]
from stereo_orb_slam_b200 import Solver, pack_window, synthetic
from stereo_orb_slam_b200.parallel import shard_points
smap = synthetic.make_config(cfg)
b = pack_window(smap, 0, smap.num_frames)
nc, npt = b["poses"].shape[0], b["points"].shape[0]
for shard in (None, 8):
    sv = Solver(0)
    if shard:
        cut = shard_points(b["obs_pt"], npt, shard)
        sv.set_point_shard(int(cut[0]), int(cut[1]))
    for rep in range(2):
        t = time.perf_counter()
        sv.set_structure(nc, npt, b["obs_cam"], b["obs_pt"])
        print("[struct_prof] %s shard %s rep %d: set_structure %.1f ms" % (cfg, shard, rep, 1e3 * (time.perf_counter() - t)), file=sys.stderr, flush=True)
    sv.close()
