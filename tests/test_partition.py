"""Multi-GPU host logic on CPU: worlds partitioned by index (no collective on the step path), per-world seed streams
independent of the rank count, and the optional episode-statistics all-reduce over a world_size-2 gloo group."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import model_path


def test_world_ranges_tile_the_batch():
    from myosuite_mjx_b200 import partition
    n = 8192
    for ws in (1, 2, 4, 8):
        seen = []
        for r in range(ws):
            seen += list(partition.world_range(r, ws, n))
        assert seen == list(range(ws * n))
    with pytest.raises(ValueError):
        partition.world_range(2, 2, 8)


def test_seeds_do_not_depend_on_rank_count(models):
    from myosuite_mjx_b200 import partition
    m = models("hand")
    whole = partition.initial_qpos(m, range(16), 1234)
    halves = np.concatenate([partition.initial_qpos(m, partition.world_range(r, 2, 8), 1234) for r in range(2)])
    assert np.array_equal(whole, halves)
    lo, hi = np.asarray(m.jnt_range)[:, 0], np.asarray(m.jnt_range)[:, 1]
    assert np.all(whole >= lo - 1e-12) and np.all(whole <= hi + 1e-12)
    assert np.abs(whole - whole[0]).max() > 0
    c = partition.random_ctrl(m.nu, range(4, 8), 1234, 3)
    assert np.array_equal(c, partition.random_ctrl(m.nu, range(0, 8), 1234, 3)[4:])
    assert c.min() >= 0 and c.max() <= 1


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from myosuite_mjx_b200 import mjcf, partition
    m = mjcf.Model.load(model_path("elbow"))
    worlds = partition.world_range(rank, world, 6)
    q0 = partition.initial_qpos(m, worlds, 99)
    status = torch.tensor([1 if (w % 4 == 0) else 0 for w in worlds], dtype=torch.int32)
    qvel = torch.full((6, 1), float(rank + 1))
    stats = partition.all_reduce_stats(partition.episode_stats(status, qvel, reward=torch.ones(6)))
    q.put((rank, stats.tolist(), q0.tolist()))
    dist.barrier()
    dist.destroy_process_group()


def test_gloo_two_ranks_stats_allreduce(models):
    from myosuite_mjx_b200 import partition
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    # both ranks hold the same global sums: 12 worlds, worlds 0,4,8 flagged, sum|qvel| = 6*1 + 6*2, reward 12
    for _, stats, _ in res:
        assert stats == [12.0, 3.0, 18.0, 12.0]
    whole = partition.initial_qpos(models("elbow"), range(12), 99)
    assert np.array_equal(np.concatenate([np.array(res[0][2]), np.array(res[1][2])]), whole)
