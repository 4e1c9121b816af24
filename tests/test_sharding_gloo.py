"""world_size-2 gloo run of the multi-GPU plumbing (env sharding, transition gather, action scatter,
episode-stat reduction) on CPU tensors: the same code path that runs over NCCL on the GPU box."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from gripper_mujoco_b200.sharding import (TransitionGather, owner_of, reduce_episode_stats, scatter_actions,
                                          shard_range)


def test_shard_range_partitions():
    for n, w in [(4096, 8), (10, 3), (7, 8), (131072, 8), (5, 1)]:
        shards = [shard_range(n, r, w) for r in range(w)]
        assert shards[0].start == 0 and shards[-1].stop == n
        assert all(a.stop == b.start for a, b in zip(shards, shards[1:]))
        assert max(s.count for s in shards) - min(s.count for s in shards) <= 1
        for e in range(0, n, max(1, n // 50)):
            r = owner_of(e, n, w)
            assert shards[r].start <= e < shards[r].stop


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _worker(rank, world, port, n_total, n_obs, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sh = shard_range(n_total, rank, world)
    ids = torch.arange(sh.start, sh.stop, dtype=torch.float32)
    # synthetic per-env transition whose content encodes the global env id
    obs = ids[:, None] * 0.001 + torch.arange(n_obs, dtype=torch.float32)[None, :]
    rew = -ids
    done = (torch.arange(sh.start, sh.stop) % 3 == 0).to(torch.uint8)
    g = TransitionGather(sh, n_obs, "cpu")
    res = g.gather(obs, rew, done)
    full_actions = torch.arange(n_total, dtype=torch.int32) % 15 if rank == 0 else None
    mine = scatter_actions(full_actions, sh, "cpu")
    stats = reduce_episode_stats(torch.tensor([float(sh.count), float(rank + 1)]))
    ok = bool((mine == (torch.arange(sh.start, sh.stop) % 15).to(torch.int32)).all())
    if rank == 0:
        o, r, d = res
        ids_all = torch.arange(n_total, dtype=torch.float32)
        ok &= o.shape == (n_total, n_obs) and bool(torch.equal(o[:, 0], ids_all * 0.001))
        ok &= bool(torch.equal(r, -ids_all)) and bool(torch.equal(d, torch.arange(n_total) % 3 == 0))
        ok &= stats.tolist() == [float(n_total), float(sum(range(1, world + 1)))]
    else:
        ok &= res is None
    np.save(os.path.join(out_dir, f"ok{rank}.npy"), np.array([ok]))
    dist.destroy_process_group()


@pytest.mark.parametrize("n_total", [64, 37])
def test_gather_scatter_world2(tmp_path, n_total):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), n_total, 9, str(tmp_path)), nprocs=world, join=True)
    for r in range(world):
        assert np.load(tmp_path / f"ok{r}.npy")[0]
