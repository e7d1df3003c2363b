"""Multi-GPU host logic; the reference is single-process, so nothing here has a counterpart in it.

Two ways the path shards over the GPUs of one box, one process per GPU:

* independent local-BA windows (config 3): `shard_windows` gives each rank a contiguous run of windows; every
  rank runs the batched single-GPU solver on its run; no data-path collective.
* one large problem (configs 4-5): `shard_points` cuts the point range into contiguous runs balanced by
  observation count; a rank owns its points and ALL their observations (V, gp, W, the per-point Schur
  contributions and the back-substitution stay local), cameras are replicated, and libsba.so all-reduces the
  camera-side sums (U, gc, S, rhs, step scalars) with NCCL over NVLink inside `sba_run`.

`ShardedSolver` wires a `Solver` to a torch.distributed process group: rank 0 makes the NCCL id
(`sba_comm_unique_id`), the group broadcasts its 128 bytes (gloo or nccl, whatever the group runs on), every
rank calls `sba_comm_init` + `sba_set_point_shard`.  torch is plumbing only; no torch type crosses the C-ABI.
"""
import numpy as np

from .solver import Solver


def shard_windows(num_windows, world_size):
    """Contiguous window ranges [lo, hi) per rank, sizes differing by at most one."""
    num_windows, world_size = int(num_windows), int(world_size)
    if world_size < 1 or num_windows < 0:
        raise ValueError("bad arguments")
    base, extra = divmod(num_windows, world_size)
    out, lo = [], 0
    for r in range(world_size):
        hi = lo + base + (1 if r < extra else 0)
        out.append((lo, hi))
        lo = hi
    return out


def shard_points(obs_pt, num_points, world_size):
    """Point boundaries b[0..world] (b[0]=0, b[world]=num_points): rank r owns points [b[r], b[r+1]) and all
    their observations; cuts balance the observation count (the unit of work of K1-K3)."""
    obs_pt = np.asarray(obs_pt)
    num_points, world_size = int(num_points), int(world_size)
    if world_size < 1 or num_points < 0:
        raise ValueError("bad arguments")
    if obs_pt.size and (obs_pt.min() < 0 or obs_pt.max() >= num_points):
        raise ValueError("obs_pt out of range")
    cum = np.concatenate([[0], np.cumsum(np.bincount(obs_pt, minlength=num_points))])
    total = int(cum[-1])
    b = [0]
    for r in range(1, world_size):
        cut = int(np.searchsorted(cum, total * r / world_size, side="left"))
        b.append(min(max(cut, b[-1]), num_points))
    b.append(num_points)
    return np.asarray(b, np.int64)


def slice_windows(lo, hi, win_cam_off, win_pt_off, obs_cam, obs_pt):
    """Sub-problem made of windows [lo, hi) of a batched problem: returns (cam range, point range, obs mask,
    local win_cam_off, local win_pt_off).  Observations keep their order (the reference's scan order)."""
    win_cam_off, win_pt_off = np.asarray(win_cam_off), np.asarray(win_pt_off)
    c0, c1, p0, p1 = int(win_cam_off[lo]), int(win_cam_off[hi]), int(win_pt_off[lo]), int(win_pt_off[hi])
    obs_cam = np.asarray(obs_cam)
    mask = (obs_cam >= c0) & (obs_cam < c1)
    return (c0, c1), (p0, p1), mask, win_cam_off[lo:hi + 1] - c0, win_pt_off[lo:hi + 1] - p0


class ShardedSolver:
    """One rank of a point-sharded global BA.  `group` is a torch.distributed process group (or None for the
    default group); device is this rank's CUDA ordinal."""

    def __init__(self, device, group=None):
        import torch
        import torch.distributed as dist
        self.dist, self.group = dist, group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        self.solver = Solver(device)
        if self.world > 1:
            uid = [Solver.comm_unique_id() if self.rank == 0 else None]
            dist.broadcast_object_list(uid, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
            self.solver.comm_init(self.rank, self.world, uid[0])
        self._torch = torch

    def solve(self, poses, points, obs_cam, obs_pt, obs_uv, p_l, p_r, cam_fixed=None, options=None):
        """Every rank passes the WHOLE problem (cameras are replicated, the library keeps only the rank's
        observations on the device); returns the full solution on every rank."""
        b = shard_points(obs_pt, np.asarray(points).size // 3, self.world)
        if self.world > 1:
            self.solver.set_point_shard(int(b[self.rank]), int(b[self.rank + 1]))
        return self.solver.solve(poses, points, obs_cam, obs_pt, obs_uv, p_l, p_r, cam_fixed=cam_fixed, options=options)
