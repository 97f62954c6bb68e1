"""Multi-GPU layout of the batched step: worlds are partitioned by index, one process per GPU, and the step path has
NO collective.  The only exchange is the optional all-reduce of a small episode-statistics vector (the counterpart
of brax's metric `pmean` implied by `max_devices_per_host=8`, reference: myosuite/mjx/ppo_continuous_action.py:59).

World `w` (global index) always draws from seed stream `seed + w`, so results do not depend on the GPU count.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.distributed as dist


def world_range(rank: int, world_size: int, nworld_per_gpu: int) -> range:
    """Global world indices owned by `rank`: [rank * n, (rank + 1) * n)."""
    if not (0 <= rank < world_size) or nworld_per_gpu <= 0:
        raise ValueError("bad rank / world_size / nworld_per_gpu")
    return range(rank * nworld_per_gpu, (rank + 1) * nworld_per_gpu)


def initial_qpos(m, worlds, seed: int, spread: float = 0.1) -> np.ndarray:
    """qpos0 + U(-spread, spread) * joint range, clipped to the range, drawn per GLOBAL world index."""
    worlds = list(worlds)
    q = np.tile(np.asarray(m.qpos0, dtype=np.float64), (len(worlds), 1))
    lim = [j for j in range(m.njnt) if m.jnt_type[j] in (2, 3) and m.jnt_limited[j]]
    for k, w in enumerate(worlds):
        u = np.random.default_rng([seed, w]).uniform(-spread, spread, len(lim))
        for x, j in zip(u, lim):
            a = m.jnt_qposadr[j]
            lo, hi = m.jnt_range[j]
            q[k, a] = np.clip(q[k, a] + x * (hi - lo), lo, hi)
    return q


def random_ctrl(nu: int, worlds, seed: int, block: int) -> np.ndarray:
    """ctrl ~ U[0,1]^nu for control block `block`, per GLOBAL world index."""
    return np.stack([np.random.default_rng([seed, w, block + 1]).uniform(0.0, 1.0, nu) for w in worlds]).astype(np.float32)


_M1, _M2 = 0xBF58476D1CE4E5B9, 0x94D049BB133111EB


def _mix_np(x: np.ndarray) -> np.ndarray:
    """splitmix64 finaliser on uint64 arrays (wrapping arithmetic)"""
    x = x.astype(np.uint64)
    x ^= x >> np.uint64(30); x *= np.uint64(_M1)
    x ^= x >> np.uint64(27); x *= np.uint64(_M2)
    x ^= x >> np.uint64(31)
    return x


def ctrl_hash_np(nu: int, worlds, seed: int, block: int) -> np.ndarray:
    """ctrl ~ U[0,1)^nu of control block `block`, a counter-based hash of (seed, GLOBAL world index, block, actuator):
    nothing to store, identical on every partition of the batch.  Host twin of `ctrl_hash_torch`."""
    w = np.asarray(list(worlds), dtype=np.uint64)[:, None]
    a = np.arange(nu, dtype=np.uint64)[None, :]
    with np.errstate(over="ignore"):
        key = _mix_np(_mix_np(np.uint64(seed) * np.uint64(0x9E3779B97F4A7C15) + w) + np.uint64(block) * np.uint64(0xD1B54A32D192ED03) + a)
    return ((key >> np.uint64(40)).astype(np.float32) * np.float32(1.0 / (1 << 24))).astype(np.float32)


def _mix_torch(x: torch.Tensor) -> torch.Tensor:
    def shr(v, n):   # logical shift right on int64
        return (v >> n) & ((1 << (64 - n)) - 1)
    c1 = _M1 - (1 << 64); c2 = _M2 - (1 << 64)   # the constants as signed int64
    x = (x ^ shr(x, 30)) * c1
    x = (x ^ shr(x, 27)) * c2
    return x ^ shr(x, 31)


def ctrl_hash_torch(nu: int, w0: int, nworld: int, seed: int, block: int, device) -> torch.Tensor:
    """Device twin of `ctrl_hash_np` for the worlds [w0, w0 + nworld): returns the storage layout [nu, nworld] float32."""
    w = torch.arange(w0, w0 + nworld, dtype=torch.int64, device=device)[None, :]
    a = torch.arange(nu, dtype=torch.int64, device=device)[:, None]
    def w64(x):   # wrap a Python int to signed 64 bits
        return ((x + (1 << 63)) % (1 << 64)) - (1 << 63)
    key = _mix_torch(_mix_torch(w64(seed * 0x9E3779B97F4A7C15) + w) + w64(block * 0xD1B54A32D192ED03) + a)
    return (((key >> 40) & ((1 << 24) - 1)).to(torch.float32) * (1.0 / (1 << 24))).contiguous()


def episode_stats(status: torch.Tensor, qvel: torch.Tensor, reward: torch.Tensor | None = None) -> torch.Tensor:
    """Local statistics vector [worlds, flagged worlds, sum |qvel|, sum reward] (float64 for exact integer sums)."""
    r = reward.double().sum() if reward is not None else torch.zeros((), dtype=torch.float64, device=status.device)
    return torch.stack([torch.tensor(float(status.numel()), dtype=torch.float64, device=status.device),
                        status.ne(0).sum().double(), qvel.abs().double().sum(), r])


def all_reduce_stats(stats: torch.Tensor) -> torch.Tensor:
    """Sum over ranks (NCCL on GPUs, gloo in the CPU tests); identity when not initialised."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(stats, op=dist.ReduceOp.SUM)
    return stats
