"""CPU, world_size 2 over gloo: the env-range sharding used by bench.py / multi-GPU rollouts.  Each rank owns a
contiguous env range; the shards' reset states (Philox subsequence = global env index) must equal the matching
slice of the single-process job, and the episode-statistics reduce must sum over ranks."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from gym_pybullet_adrp_b200 import _abi, configs
from gym_pybullet_adrp_b200.sharding import reduce_episode_stats, shard_range
from gym_pybullet_adrp_b200.utils.enums import RaceMode


def test_shard_range_partitions():
    for total in (1, 7, 8, 1000, 131072):
        for world in (1, 2, 3, 4, 8):
            spans = [shard_range(total, world, r) for r in range(world)]
            assert spans[0][0] == 0 and sum(c for _, c in spans) == total
            for (o0, c0), (o1, _) in zip(spans, spans[1:]):
                assert o0 + c0 == o1
            assert max(c for _, c in spans) - min(c for _, c in spans) <= 1
    with pytest.raises(ValueError):
        shard_range(8, 2, 2)


def _race_cfg(n, offset):
    return _abi.make_race_config(configs.load_config("level3"), num_envs=n, num_drones=2, racemode=RaceMode.COMPETE,
                                 precision=_abi.F64, seed=1337, env_offset=offset)


def _worker(rank, world, port, total, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from tests.oracle_lib import OracleEnv
        off, cnt = shard_range(total, world, rank)
        env = OracleEnv(_race_cfg(cnt, off))               # the CPU checker stands in for the GPU handle
        a = np.zeros((cnt, 2, 4)); a[..., 0:2] = [[0.9, 0.9], [1.1, 1.1]]; a[..., 2] = 0.4
        for _ in range(3):
            obs, rew, te, tr, fl = env.step(a)
        stats = torch.tensor([float(obs[..., 2].sum()), float(cnt), float((fl & _abi.FLAG_ELIMINATED).astype(bool).sum())],
                             dtype=torch.float64)
        reduce_episode_stats(stats)
        np.save(os.path.join(out, f"state_{rank}.npy"), env.get_state())
        np.save(os.path.join(out, f"stats_{rank}.npy"), stats.numpy())
        dist.barrier()
    finally:
        dist.destroy_process_group()


def test_two_rank_shards_match_single_process(tmp_path):
    from tests.oracle_lib import OracleEnv
    total, world = 37, 2
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mp.spawn(_worker, args=(world, port, total, str(tmp_path)), nprocs=world, join=True)
    whole = OracleEnv(_race_cfg(total, 0))
    a = np.zeros((total, 2, 4)); a[..., 0:2] = [[0.9, 0.9], [1.1, 1.1]]; a[..., 2] = 0.4
    for _ in range(3):
        obs, rew, te, tr, fl = whole.step(a)
    parts = [np.load(tmp_path / f"state_{r}.npy") for r in range(world)]
    np.testing.assert_array_equal(np.concatenate(parts, axis=0), whole.get_state())      # independent of the world size
    want = np.array([obs[..., 2].sum(), total, (fl & _abi.FLAG_ELIMINATED).astype(bool).sum()], dtype=np.float64)
    for r in range(world):
        np.testing.assert_allclose(np.load(tmp_path / f"stats_{r}.npy"), want, rtol=1e-12)
