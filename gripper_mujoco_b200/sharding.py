"""Multi-GPU plumbing: environments shard by index, one process per GPU, and the only exchange is the
gather of transitions / episode statistics to the learner rank. There is no collective
inside the physics: the reference runs every MjClass in its own OS process (rl/Trainer.py), and the
counter RNG is keyed by the GLOBAL env id, so results do not depend on the number of GPUs.

Works over any torch.distributed backend: "nccl" (NVLink / NVSwitch) on the B200 box, "gloo" in the
CPU tests (tests/test_sharding_gloo.py).
"""
from __future__ import annotations

from dataclasses import dataclass

import torch
import torch.distributed as dist


@dataclass
class Shard:
    rank: int
    world: int
    n_total: int
    start: int
    count: int

    @property
    def stop(self) -> int:
        return self.start + self.count


def shard_range(n_total: int, rank: int, world: int) -> Shard:
    """Contiguous block partition, remainder spread over the first ranks (env e -> rank e // per)."""
    base, rem = divmod(n_total, world)
    count = base + (1 if rank < rem else 0)
    start = rank * base + min(rank, rem)
    return Shard(rank, world, n_total, start, count)


def owner_of(env: int, n_total: int, world: int) -> int:
    base, rem = divmod(n_total, world)
    edge = rem * (base + 1)
    return env // (base + 1) if env < edge else rem + (env - edge) // max(base, 1)


class TransitionGather:
    """Gathers (next_obs, reward, done) of every shard to `dst` once per action-step.

    The per-rank payload is packed into ONE contiguous float32 buffer [count, n_obs + 2] so the exchange
    is a single collective per step (launch latency, not bandwidth, is what matters at these sizes:
    16384 envs x 86 floats = 5.6 MB per GPU). Uneven shards are padded to the largest shard.
    """

    def __init__(self, shard: Shard, n_obs: int, device, dst: int = 0, group=None):
        self.shard, self.n_obs, self.dst, self.group = shard, n_obs, dst, group
        self.width = n_obs + 2
        self.max_count = -(-shard.n_total // shard.world)
        self.send = torch.zeros((self.max_count, self.width), dtype=torch.float32, device=device)
        self.recv = None
        if shard.rank == dst:
            self.recv = [torch.zeros_like(self.send) for _ in range(shard.world)]

    def pack(self, obs: torch.Tensor, reward: torch.Tensor, done: torch.Tensor) -> torch.Tensor:
        c = self.shard.count
        self.send[:c, :self.n_obs].copy_(obs)
        self.send[:c, self.n_obs].copy_(reward)
        self.send[:c, self.n_obs + 1].copy_(done.to(torch.float32))
        return self.send

    def gather(self, obs, reward, done):
        """Returns (obs [n_total, n_obs], reward [n_total], done [n_total] bool) on dst, None elsewhere."""
        self.pack(obs, reward, done)
        if self.shard.world == 1:
            full = self.send[: self.shard.count]
            return full[:, :self.n_obs], full[:, self.n_obs], full[:, self.n_obs + 1] > 0.5
        dist.gather(self.send, self.recv if self.shard.rank == self.dst else None, dst=self.dst, group=self.group)
        if self.shard.rank != self.dst:
            return None
        parts = [self.recv[r][: shard_range(self.shard.n_total, r, self.shard.world).count] for r in range(self.shard.world)]
        full = torch.cat(parts, dim=0)
        return full[:, :self.n_obs], full[:, self.n_obs], full[:, self.n_obs + 1] > 0.5


def scatter_actions(actions_full, shard: Shard, device, src: int = 0, group=None) -> torch.Tensor:
    """Learner rank holds int32 actions for all envs; every rank receives its slice."""
    max_count = -(-shard.n_total // shard.world)
    out = torch.zeros(max_count, dtype=torch.int32, device=device)
    if shard.world == 1:
        return actions_full[shard.start:shard.stop].to(device=device, dtype=torch.int32)
    chunks = None
    if shard.rank == src:
        chunks = []
        for r in range(shard.world):
            s = shard_range(shard.n_total, r, shard.world)
            c = torch.zeros(max_count, dtype=torch.int32, device=device)
            c[: s.count].copy_(actions_full[s.start:s.stop])
            chunks.append(c)
    dist.scatter(out, chunks, src=src, group=group)
    return out[: shard.count]


def reduce_episode_stats(stats: torch.Tensor, dst: int = 0, group=None) -> torch.Tensor:
    """Sum of per-rank episode statistics (returns, lengths, event counts) on dst."""
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.reduce(stats, dst=dst, op=dist.ReduceOp.SUM, group=group)
    return stats
