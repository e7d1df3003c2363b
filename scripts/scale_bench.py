#!/usr/bin/env python
"""Multi-GPU measurements for the two ways the path shards; bench.py stays the single contract line.

  torchrun --nproc-per-node N scripts/scale_bench.py --mode C3 [--windows 4096]   independent local-BA windows,
      windows dealt to the ranks (shard_windows), no data-path collective            -> strong scaling over N
  torchrun --nproc-per-node N scripts/scale_bench.py --mode C4 [--scale 1.0]      one global BA, points sharded
      (shard_points), camera-side sums all-reduced by NCCL inside libsba.so          -> strong scaling over N
  ... --mode C5                                                                   the 20k-pose / 5M-point / 30M-observation stress config, same path as C4

Each prints one JSON line on rank 0: K1-K3 observations/s (device time, max over ranks, L2 flushed between
steps) and LM iterations/s of a capped solve (sba_run, state resident), plus the final cost for the parity table.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def batched_problem(nwin, base=16):
    """nwin windows of config C3 (10 keyframes / 5k points), `base` distinct ones tiled; flat batched arrays."""
    from stereo_orb_slam_b200 import pack_window, synthetic
    ws = []
    for i in range(min(base, nwin)):
        smap = synthetic.make_config("C3", seed=synthetic.CONFIGS["C3"]["seed"] + i)
        w = pack_window(smap, 0, smap.num_frames); w["p_l"], w["p_r"] = smap.p_l, smap.p_r
        ws.append(w)
    pick = [ws[i % len(ws)] for i in range(nwin)]
    cam_off = np.cumsum([0] + [w["poses"].shape[0] for w in pick]).astype(np.int32)
    pt_off = np.cumsum([0] + [w["points"].shape[0] for w in pick]).astype(np.int32)
    return dict(poses=np.concatenate([w["poses"] for w in pick]), points=np.concatenate([w["points"] for w in pick]),
                obs_cam=np.concatenate([w["obs_cam"] + cam_off[i] for i, w in enumerate(pick)]).astype(np.int32),
                obs_pt=np.concatenate([w["obs_pt"] + pt_off[i] for i, w in enumerate(pick)]).astype(np.int32),
                obs_uv=np.concatenate([w["obs_uv"] for w in pick]), win_cam_off=cam_off, win_pt_off=pt_off,
                p_l=ws[0]["p_l"], p_r=ws[0]["p_r"])


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mode", default="C3", choices=["C3", "C4", "C5"])
    ap.add_argument("--windows", type=int, default=4096)
    ap.add_argument("--scale", type=float, default=1.0)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--iters", type=int, default=10, help="LM iteration cap of the timed solve")
    args = ap.parse_args()
    sys.stdout.flush(); json_out = os.fdopen(os.dup(1), "w"); os.dup2(2, 1)     # stdout carries the JSON line only (NCCL's banner goes to stderr)
    import torch
    import torch.distributed as dist
    from stereo_orb_slam_b200 import Solver, default_options, pack_window, synthetic
    from stereo_orb_slam_b200.parallel import shard_points, shard_windows, slice_windows

    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")     # stdout carries the JSON line only
        os.environ.setdefault("SBA_HOST_THREADS", str(max(1, (os.cpu_count() or 1) // world)))
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    opt = default_options(max_time_s=0.0, max_iterations=args.iters, verbose=1 if os.environ.get("SCALE_VERBOSE") else 0)
    sv = Solver(local)
    if args.mode == "C3":
        lo, hi = shard_windows(args.windows, world)[rank]
        b = batched_problem(hi - lo)             # every rank builds only its own windows
        sv.set_structure(b["poses"].shape[0], b["points"].shape[0], b["obs_cam"], b["obs_pt"], win_cam_off=b["win_cam_off"],
                         win_pt_off=b["win_pt_off"])
        nobs_local, label = int(b["obs_cam"].size), "C3: %d independent windows of 10 keyframes / 5k points" % args.windows
    else:
        smap = synthetic.make_config(args.mode, scale=args.scale)
        b = pack_window(smap, 0, smap.num_frames); b["p_l"], b["p_r"] = smap.p_l, smap.p_r
        if world > 1:
            uid = [Solver.comm_unique_id() if rank == 0 else None]
            dist.broadcast_object_list(uid, src=0)
            sv.comm_init(rank, world, uid[0])
            cut = shard_points(b["obs_pt"], b["points"].shape[0], world)
            sv.set_point_shard(int(cut[rank]), int(cut[rank + 1]))
            nobs_local = int(np.count_nonzero((b["obs_pt"] >= cut[rank]) & (b["obs_pt"] < cut[rank + 1])))
        else:
            nobs_local = int(b["obs_cam"].size)
        sv.set_structure(b["poses"].shape[0], b["points"].shape[0], b["obs_cam"], b["obs_pt"])
        label = "%s x%.2f: %d poses / %d points / %d observations, points sharded" % (args.mode, args.scale, b["poses"].shape[0], b["points"].shape[0], b["obs_cam"].size)
    sv.set_state(b["poses"], b["points"]); sv.set_measurements(b["obs_uv"], b["p_l"], b["p_r"])
    sv.linearize()
    for _ in range(3):
        sv.time_stage(8, 1, True, 1e4)
    barrier()
    ms = sum(sv.time_stage(8, 1, True, 1e4) for _ in range(args.steps))
    barrier()
    # the solve kernels (factorisation, step loop) load and set up on their first launch: two untimed LM iterations first
    sv.set_state(b["poses"], b["points"])
    sv.run(default_options(max_time_s=0.0, max_iterations=2))
    sv.set_state(b["poses"], b["points"])
    barrier(); t0 = time.perf_counter()
    s = sv.run(opt)
    barrier(); wall = time.perf_counter() - t0
    t = torch.tensor([ms, wall], dtype=torch.float64, device="cuda"); c = torch.tensor([float(nobs_local), s["final_cost"] if args.mode == "C3" else 0.0], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX); dist.all_reduce(c, op=dist.ReduceOp.SUM)
    if rank == 0:
        nobs = float(c[0])
        final_cost = float(c[1]) if args.mode == "C3" else s["final_cost"]
        json_out.write(json.dumps({"mode": args.mode, "workload": label, "n_gpus": world, "scaling": "strong",
                          "k1k3_obs_per_s": nobs * args.steps / (float(t[0]) * 1e-3), "ms_per_step": float(t[0]) / args.steps,
                          "lm_iterations": s["iterations"], "lm_iters_per_s": s["iterations"] / float(t[1]), "solve_s": float(t[1]),
                          "pcg_iterations": s["pcg_iterations"], "final_cost": final_cost, "termination": s["termination_name"],
                          "device_ms_rank0": {k: s[k] for k in ("ms_linearize", "ms_schur", "ms_pcg", "ms_update")}}) + "\n"); json_out.flush()
    sv.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
