"""Multi-GPU host logic: shard planning (pure), world_size-2 gloo runs of the structure build on host-only
handles (CPU), and — on a box with >= 2 GPUs — the NCCL point-sharded solve against the single-GPU solve."""
import os
import socket

import numpy as np
import pytest

from stereo_orb_slam_b200.parallel import shard_points, shard_windows

import _mp_workers as workers


def free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def spawn(fn, world, out_dir):
    import torch.multiprocessing as mp
    mp.spawn(fn, args=(world, free_port(), str(out_dir)), nprocs=world, join=True)


def test_shard_windows_tiles_the_batch():
    assert shard_windows(4096, 8) == [(512 * r, 512 * (r + 1)) for r in range(8)]
    assert shard_windows(5, 2) == [(0, 3), (3, 5)]
    assert shard_windows(1, 4) == [(0, 1), (1, 1), (1, 1), (1, 1)]
    assert shard_windows(0, 2) == [(0, 0), (0, 0)]


def test_shard_points_balances_observations():
    rng = np.random.default_rng(5)
    counts = rng.integers(1, 9, size=5000)
    obs_pt = np.repeat(np.arange(5000), counts)
    rng.shuffle(obs_pt)
    for world in (1, 2, 4, 8):
        b = shard_points(obs_pt, 5000, world)
        assert b[0] == 0 and b[-1] == 5000 and np.all(np.diff(b) >= 0) and b.size == world + 1
        per = [int(counts[b[r]:b[r + 1]].sum()) for r in range(world)]
        assert sum(per) == obs_pt.size and max(per) - min(per) <= 16
    assert shard_points(np.zeros(0, np.int32), 0, 4).tolist() == [0, 0, 0, 0, 0]
    with pytest.raises(ValueError):
        shard_points(np.array([7]), 5, 2)


def test_point_shards_tile_the_problem_world2_gloo(tmp_path):
    spawn(workers.host_shard_worker, 2, tmp_path)
    ok, lo, hi = (int(x) for x in open(tmp_path / "result.txt").read().split())
    assert ok == 1 and hi - lo <= 16


def test_window_shards_are_self_contained_world2_gloo(tmp_path):
    spawn(workers.window_shard_worker, 2, tmp_path)
    assert open(tmp_path / "result.txt").read().strip() == "1"


@pytest.mark.gpu
def test_point_sharded_solve_equals_single_gpu(tmp_path):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs (run with gpurun --gpus 2)")
    spawn(workers.gpu_sharded_worker, 2, tmp_path)
    r = np.load(tmp_path / "result.npz")
    assert abs(r["init"] - r["init1"]) <= 1e-12 * r["init1"]
    # north_star tolerances: cost 1e-5 relative, poses 1e-4 m / 1e-5 rad, points 1e-4 m
    assert abs(r["cost"] - r["cost1"]) <= 1e-5 * r["cost1"]
    assert np.abs(r["poses"][:, :3] - r["poses1"][:, :3]).max() < 1e-5
    assert np.abs(r["poses"][:, 3:] - r["poses1"][:, 3:]).max() < 1e-4
    assert np.abs(r["points"] - r["points1"]).max() < 1e-4


@pytest.mark.gpu
def test_point_sharded_global_problem_factors_exactly_and_times_out_together(tmp_path):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs (run with gpurun --gpus 2)")
    spawn(workers.gpu_sharded_global_worker, 2, tmp_path)
    r = np.load(tmp_path / "result.npz")
    # exact reduced solve on both sides: the sharded trajectory equals the single-GPU one to summation order
    assert int(r["pcg"]) == 0 and int(r["it"]) == int(r["it1"]) == 12
    assert abs(r["cost"] - r["cost1"]) <= 1e-9 * r["cost1"]
    assert np.abs(r["poses"] - r["poses1"]).max() < 1e-7 and np.abs(r["points"] - r["points1"]).max() < 1e-6
    # wall-clock limit: every rank stops with max_time after the SAME iteration with the same cost
    assert all(t == "max_time" for t in r["t_term"])
    assert len(set(r["t_it"].tolist())) == 1 and len(set(r["t_cost"].tolist())) == 1
    assert 0 < r["t_it"][0] < 50 and r["t_cost"][0] < r["t_init"]
