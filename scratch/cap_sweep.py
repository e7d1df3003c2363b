import sys, os, subprocess
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if len(sys.argv) > 1 and sys.argv[1] == "one":
    sys.path.insert(0, ROOT)
    from stereo_orb_slam_b200 import _capi
    if os.environ.get("SBA_LIB"): _capi.LIB_PATH = os.path.abspath(os.environ["SBA_LIB"])
    from stereo_orb_slam_b200 import Solver, pack_window, synthetic
    smap = synthetic.make_config("C2"); w = pack_window(smap, 0, smap.num_frames)
    sv = Solver(0).set_structure(w["poses"].shape[0], w["points"].shape[0], w["obs_cam"], w["obs_pt"])
    sv.set_state(w["poses"], w["points"]); sv.set_measurements(w["obs_uv"], smap.p_l, smap.p_r)
    sv.linearize(); sv.schur(1e4)
    for _ in range(2): sv.time_stage(8, 20, True, 1e4)
    r = {k: 1e3 * min(sv.time_stage(k, 20, True, 1e4) for _ in range(3)) for k in (0, 9, 2, 1, 8)}
    print(os.path.basename(os.environ.get("SBA_LIB","default")), "cap %s chunks %d  lin %.1f asm_c %.1f schur %.1f asm_s %.1f step %.1f" % (os.environ.get("SBA_OBS_CAP"), -1, r[0], r[9], r[2], r[1], r[8]), flush=True)
else:
    for cap in sys.argv[1:]:
        subprocess.run([sys.executable, __file__, "one"], env=dict(os.environ, SBA_OBS_CAP=cap))
