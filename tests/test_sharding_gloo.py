"""Host-side multi-GPU logic (parallax_b200/sharding.py) on CPU: world_size-2 `gloo` processes.
Envs are independent, so sharding the batch must not change any env's trajectory, and the only
collective is the episode-return reduction.  The per-shard step here is the CPU oracle standing
in for the kernel (this test is about the sharding logic, not the kernel)."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_shard_range_partitions_the_batch():
    from parallax_b200.sharding import shard_keys, shard_range

    for total in (0, 1, 7, 8, 1000, 1 << 20):
        for world in (1, 2, 3, 4, 8):
            spans = [shard_range(total, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == total
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard_range(10, 2, 2)
    # keys are a function of the GLOBAL env index only
    k = np.concatenate([shard_keys(10, r, 3) for r in range(3)])
    assert np.array_equal(k, shard_keys(10, 0, 1))
    assert np.array_equal(k[:, 1], np.arange(10))


def _worker(rank, world, port, total, steps, out_dir):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank),
                      WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    import torch
    import torch.distributed as dist

    from oracle import oracle as o
    from parallax_b200 import sharding as sh

    dist.init_process_group("gloo", rank=rank, world_size=world)
    assert sh.dist_env() == (rank, world, rank)
    P = o.default_lunar_params()
    P.judge_enabled, P.auto_reset, P.reset_perturb, P.max_episode_steps = 1, 1, 1, 25
    lo, hi = sh.shard_range(total, rank, world)
    st = o.lunar_reset(sh.shard_keys(total, rank, world), P)
    red = sh.ReturnReducer()
    acts = np.random.default_rng(9).uniform(-1, 1, size=(steps, 2, total)).astype(np.float32)
    for t in range(steps):
        prev_ret = st[34].copy()
        st, rew, done = o.lunar_step(st, acts[t][:, lo:hi], P)
        red.update(torch.from_numpy(prev_ret + rew), torch.from_numpy(done))
    stats = red.reduce()
    tmax = sh.max_over_ranks(float(rank + 1))
    np.save(os.path.join(out_dir, f"state_{rank}.npy"), st)
    if rank == 0:
        np.save(os.path.join(out_dir, "stats.npy"),
                np.array([stats.return_sum, stats.episodes, stats.env_steps, tmax]))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_sharding_matches_unsharded(tmp_path, oracle):
    import torch
    import torch.multiprocessing as mp

    total, steps, world = 37, 60, 2
    mp.spawn(_worker, args=(world, _free_port(), total, steps, str(tmp_path)), nprocs=world, join=True)
    got = np.concatenate([np.load(tmp_path / f"state_{r}.npy") for r in range(world)], axis=1)
    # unsharded run of the same job
    from parallax_b200 import sharding as sh

    P = oracle.default_lunar_params()
    P.judge_enabled, P.auto_reset, P.reset_perturb, P.max_episode_steps = 1, 1, 1, 25
    st = oracle.lunar_reset(sh.shard_keys(total, 0, 1), P)
    acts = np.random.default_rng(9).uniform(-1, 1, size=(steps, 2, total)).astype(np.float32)
    ret_sum = episodes = 0.0
    for t in range(steps):
        prev = st[34].copy()
        st, rew, done = oracle.lunar_step(st, acts[t], P)
        fin = done != 0
        ret_sum += float((prev + rew)[fin].astype(np.float64).sum())
        episodes += float(fin.sum())
    ok = (got == st) | (np.isnan(got) & np.isnan(st))
    assert ok.all()
    stats = np.load(tmp_path / "stats.npy")
    assert episodes >= total * 2  # max_episode_steps = 25 over 60 steps
    assert stats[1] == episodes and stats[2] == total * steps
    assert abs(stats[0] - ret_sum) <= 1e-6 * max(1.0, abs(ret_sum))
    assert stats[3] == world  # max over ranks
