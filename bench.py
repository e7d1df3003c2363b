#!/usr/bin/env python
"""bench.py — BA observations/sec through residual + Jacobian + Hessian accumulation + Schur (K1-K3),
plus LM iterations/sec, on BASELINE.json's configs[1] (C2: 50 keyframes / 40k points / ~200k stereo
observations per window).  One JSON line on stdout (rank 0).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload C2|C1|C4s]

ours       value = observations through K1-K3 per second, state resident in HBM, device time (CUDA events on
           the solver's stream), L2 flushed between steps; N > 1: every rank owns one independent window
           (weak scaling, no data-path collective — SURVEY §8(e)), value = all ranks' observations / max time.
           e2e   = the same unit through the C-ABI one-shot call sba_solve on HOST buffers (structure build,
                   H2D, the whole LM solve, D2H inside the timed region): observations x linearisations / wall.
reference  the CPU restatement of the reference's Ceres path (oracle/, "port": Ceres itself cannot be built
           here) on the box's host cores: the LM solve capped at a few iterations per step, same unit.
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "BA observations/sec through residual+Jacobian+Hessian accumulation+Schur (K1-K3)"
UNIT = "obs/s"
# dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed `ncu --set full` capture of the same
# workload (profiles/, see the summary there); None for a kernel without a capture
TRAFFIC = {"k_schur_chunk": 36198656, "k_lin_chunk": 6119168}   # profiles/r01_f_ncu_full_summary.txt (C2; W written by k_lin_chunk stays in L2)


def make_workload(name, rank=0):
    from stereo_orb_slam_b200 import pack_window, synthetic
    if name == "C4s":          # 1/10-scale global BA (454 poses / 100k points), single GPU
        smap = synthetic.make_config("C4", scale=0.1, seed=synthetic.CONFIGS["C4"]["seed"] + rank)
    else:
        # every rank solves a window of the same synthetic map (weak scaling: per-GPU work is fixed and identical)
        smap = synthetic.make_config(name)
    w = pack_window(smap, 0, smap.num_frames)
    w["p_l"], w["p_r"] = smap.p_l, smap.p_r
    fixed = np.zeros(smap.num_frames, np.uint8); fixed[0] = 1
    w["cam_fixed"] = fixed
    return w


def algorithmic_bytes(Nobs, Np, Nc, nblk):
    """SURVEY §8(d) compulsory HBM traffic."""
    lin = Nobs * 168 + Np * 96 + Nc * 264
    schur = Nobs * 144 + Np * 72 + nblk * 288 + Nc * 48
    return lin, schur


class ClockSampler(threading.Thread):
    """nvidia-smi style clock / throttle-reason samples during the timed region (NVML)."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.stop_flag, self.max_mhz = index, [], set(), False, None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
                 nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap"}
        while not self.stop_flag:
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, n in names.items():
                    if r & bit:
                        self.reasons.add(n)
            except Exception:
                pass
            time.sleep(0.02)

    def result(self):
        self.stop_flag = True
        if self.nv is None or not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["nvml unavailable"]}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, burst copy)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def run_reference(args):
    """The reference arm: CPU restatement of the reference's Ceres path on the host cores (rank 0 only)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import oracle
    w = make_workload(args.workload)
    Nobs = int(w["obs_cam"].size)
    cores = oracle.num_procs()
    cap = 3        # LM iterations per step: a bounded sample of the solve (same per-iteration work as the full solve)
    opt = oracle.default_options(max_time_s=0.0, num_threads=0, max_iterations=cap)

    def step():
        t0 = time.perf_counter()
        _, _, s = oracle.solve(w["poses"], w["points"], w["obs_cam"], w["obs_pt"], w["obs_uv"], w["p_l"], w["p_r"],
                               options=opt)
        return time.perf_counter() - t0, s
    for _ in range(min(args.warmup, 1)):
        step()
    tot, lin, its = 0.0, 0, 0
    for _ in range(args.steps):
        dt, s = step()
        tot += dt; lin += s["linearizations"]; its += s["iterations"]
    value = Nobs * lin / tot
    k13 = oracle.time_linearize_schur(w["poses"], w["points"], w["obs_cam"], w["obs_pt"], w["obs_uv"], w["p_l"], w["p_r"],
                                      w["cam_fixed"], num_threads=0, repeats=2)
    sample = "%d LM solves capped at %d iterations each on %s (%d obs); Jet autodiff + Schur + dense Cholesky" % (
        args.steps, cap, args.workload, Nobs)
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * tot / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "%s: %d keyframes / %d points / %d stereo observations" % (
                args.workload, w["poses"].shape[0], w["points"].shape[0], Nobs)},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "lm_iters_per_s": its / tot, "k1k3_obs_per_s": Nobs / k13,
            "note": "CPU restatement of the reference's Ceres path — Ceres itself unavailable on this box"}
    emit(line)


_JSON_OUT = None


def emit(line):
    out = _JSON_OUT or sys.stdout
    out.write(json.dumps(line) + "\n"); out.flush()


def run_ours(args):
    import torch
    import torch.distributed as dist
    from stereo_orb_slam_b200 import Solver, default_options

    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and "SBA_HOST_THREADS" not in os.environ:       # torchrun pins OMP_NUM_THREADS=1; share the host cores instead
        os.environ["SBA_HOST_THREADS"] = str(max(1, (os.cpu_count() or 1) // world))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        # keep stdout to the one JSON line: NCCL writes its banner ("NCCL version ...", printed from level VERSION
        # up, WARN included) and anything else it logs to stderr instead
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    w = make_workload(args.workload, rank)
    Nc, Np, Nobs = w["poses"].shape[0], w["points"].shape[0], int(w["obs_cam"].size)
    sv = Solver(local).set_structure(Nc, Np, w["obs_cam"], w["obs_pt"])
    sv.set_state(w["poses"], w["points"]); sv.set_measurements(w["obs_uv"], w["p_l"], w["p_r"])
    sv.linearize()
    radius = 1e4
    nblk = int(sv.schur_pattern()[0].size)
    B_lin, B_schur = algorithmic_bytes(Nobs, Np, Nc, nblk)
    opt = default_options(max_time_s=0.0)

    # ---- value: K steps of (linearise + Schur), cold L2 each step, device time on the solver's stream
    for _ in range(max(args.warmup, 3)):
        sv.time_stage(8, 1, True, radius)
    sampler = ClockSampler(local); sampler.start()
    barrier()
    l0 = sv.launch_count
    ms = 0.0
    for _ in range(args.steps):
        ms += sv.time_stage(8, 1, True, radius)
    barrier()
    launches = sv.launch_count - l0
    clocks = sampler.result()
    t = torch.tensor([ms, float(Nobs)], dtype=torch.float64, device="cuda")
    if world > 1:
        tm = t.clone(); dist.all_reduce(tm, op=dist.ReduceOp.MAX)
        ts = t.clone(); dist.all_reduce(ts, op=dist.ReduceOp.SUM)
        ms_max, nobs_all = float(tm[0]), float(ts[1])
    else:
        ms_max, nobs_all = ms, float(Nobs)
    value = nobs_all * args.steps / (ms_max * 1e-3)

    # ---- per-stage device times for the roofline (same stream, CUDA events, cold L2)
    reps = max(args.steps, 5)
    t_lp = sv.time_stage(0, reps, True, radius); t_lc = sv.time_stage(1, reps, True, radius)
    t_sp = sv.time_stage(2, reps, True, radius); t_lin = sv.time_stage(3, reps, True, radius)
    t_sch = sv.time_stage(4, reps, True, radius)
    peak, peak_src = measured_peak()
    # the dominant kernel of the step: k_lin_chunk does the whole of K1/K2 and k_schur_chunk the whole of K3 (their
    # assembly is in-kernel), so each is charged its stage's full algorithmic bytes
    if t_sp >= t_lp:
        dom, t_dom, B_dom = "k_schur_chunk", t_sp, B_schur
    else:
        dom, t_dom, B_dom = "k_lin_chunk", t_lp, B_lin
    ach = B_dom / (t_dom * 1e-3) / 1e9

    # ---- e2e: sba_solve on host buffers (pinned), whole LM solve, results read back
    def pinned(a):
        tt = torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
        return tt.numpy(), tt
    keep = []
    hb = {}
    for k in ("poses", "points", "obs_cam", "obs_pt", "obs_uv"):
        hb[k], tt = pinned(w[k]); keep.append(tt)
    sv2 = Solver(local)
    for _ in range(1):
        sv2.solve(hb["poses"], hb["points"], hb["obs_cam"], hb["obs_pt"], hb["obs_uv"], w["p_l"], w["p_r"], options=opt)
    e_steps = max(3, min(args.steps, 10))
    barrier()
    t0 = time.perf_counter(); lin = 0; its = 0; solve_s = 0.0; summ = None; brk = {}
    for _ in range(e_steps):
        _, _, summ = sv2.solve(hb["poses"], hb["points"], hb["obs_cam"], hb["obs_pt"], hb["obs_uv"], w["p_l"], w["p_r"], options=opt)
        lin += summ["linearizations"]; its += summ["iterations"]; solve_s += summ["time_solve_s"]
        brk = {k: brk.get(k, 0.0) + 1e3 * summ[k] / e_steps for k in ("time_structure_s", "time_h2d_s", "time_solve_s", "time_d2h_s")}
    barrier()
    e_wall = time.perf_counter() - t0
    te = torch.tensor([e_wall], dtype=torch.float64, device="cuda"); tl = torch.tensor([float(Nobs * lin)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX); dist.all_reduce(tl, op=dist.ReduceOp.SUM)
    e2e = float(tl[0]) / float(te[0])
    h2d = sum(int(hb[k].nbytes) for k in hb) + 192
    d2h = int(hb["poses"].nbytes + hb["points"].nbytes)

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_max / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": "%s: %d keyframes / %d points / %d stereo observations per window, one window per GPU" % (
                args.workload, Nc, Np, Nobs), "l2": "flushed between steps (256 MiB memset)", "residual_dim": 4,
                "schur_blocks": nblk},
            "clocks": clocks, "gpu_launches": int(launches),
            "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "host_threads_per_rank": int(os.environ.get("SBA_HOST_THREADS", os.environ.get("OMP_NUM_THREADS", os.cpu_count() or 1))), "what": "sba_solve on pinned host buffers: structure build + H2D + full LM solve + D2H; obs x linearisations / wall",
                    "ms_per_solve": 1e3 * e_wall / e_steps, "lm_iterations": its // e_steps, "final_cost": summ["final_cost"],
                    "termination": summ["termination_name"],
                    "ms_breakdown": {k.replace("time_", "").replace("_s", ""): v for k, v in brk.items()},
                    "device_ms": {k: summ[k] for k in ("ms_linearize", "ms_schur", "ms_pcg", "ms_update")}},
            "lm_iters_per_s": its / solve_s,
            "roofline": {"bound": "hbm", "kernel": dom, "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                         "traffic": TRAFFIC.get(dom), "peak_source": peak_src, "algorithmic_bytes": B_dom, "us_per_launch": 1e3 * t_dom},
            "stages_us": {"k_lin_chunk": 1e3 * t_lp, "k_assemble_schur": 1e3 * t_lc, "k_schur_chunk": 1e3 * t_sp, "linearize": 1e3 * t_lin,
                          "schur": 1e3 * t_sch},
            "stage_roofline": {"linearize_GBs": B_lin / (t_lin * 1e-3) / 1e9, "schur_GBs": B_schur / (t_sch * 1e-3) / 1e9,
                               "linearize_frac": B_lin / (t_lin * 1e-3) / 1e9 / peak, "schur_frac": B_schur / (t_sch * 1e-3) / 1e9 / peak}}
    if rank == 0 and world == 1 and not args.no_cpu:
        from oracle import oracle
        cores = oracle.num_procs()
        reps_cpu = 3
        sec = oracle.time_linearize_schur(w["poses"], w["points"], w["obs_cam"], w["obs_pt"], w["obs_uv"], w["p_l"], w["p_r"],
                                          w["cam_fixed"], num_threads=0, repeats=reps_cpu)
        line["cpu_baseline"] = {"value": Nobs / sec, "unit": UNIT, "cores": cores, "kind": "port",
                                "sample": "%d passes of linearise (Jet autodiff) + block build + dense Schur over the same %s window" % (reps_cpu, args.workload)}
    if rank == 0:
        emit(line)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="C2", choices=["C1", "C2", "C4s"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    args = ap.parse_args()
    # stdout carries exactly ONE JSON line: file descriptor 1 is pointed at stderr for everything else (NCCL prints
    # its version banner to stdout whatever NCCL_DEBUG_FILE says) and the line goes to the saved descriptor
    global _JSON_OUT
    sys.stdout.flush()
    _JSON_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
