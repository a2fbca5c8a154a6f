"""GPU suite, needs >= 2 visible GPUs (skips otherwise): two ranks, each stepping half of the tables on its own GPU,
reduce their statistics with the path's one collective (NCCL all-reduce of the RunStats / HandStats vectors, SURVEY.md
8(e)) and must reproduce what ONE full-size engine accumulates -- counts and the integer hand statistics bit for bit,
the float reward sums to rounding (the order of a floating-point sum differs)."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
N_TOTAL, STEPS, SEED = 1 << 18, 24, 20260401
KW = dict(n_players=2, max_players=2, starting_bbs=100, max_bbs=1000, w1=.5, w2=.3, K=100, alpha=50)


def _run(dev, n, env_base, actions, dist=None):
    import torch
    from pulselib_b200.envs.poker_packed import PokerGPUPacked
    from pulselib_b200.hand_stats import HandStats
    from pulselib_b200.sharding import RunStats

    env = PokerGPUPacked(dev, n_games=n, seed=SEED, env_base=env_base, auto_reset=True, **KW)
    rs, hs = RunStats(dev), HandStats(env, q_seat=0)
    env.reset()
    hs.begin_hand()
    a = torch.from_numpy(actions).to(dev)
    for t in range(STEPS):
        obs, rew, done, _, _ = env.step(a[t])
        rs.update(rew, done)
        hs.update(done)
    run = rs.all_reduce() if dist is not None else rs.vec.clone()
    if dist is not None:
        hs.allreduce()
    return run.cpu().numpy(), hs.vec.cpu().numpy()


def _worker(rank, world, port, q):
    import torch
    import torch.distributed as dist

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dev = torch.device("cuda", rank)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    from pulselib_b200.sharding import shard_range

    base, n = shard_range(N_TOTAL, rank, world)
    acts = np.random.default_rng(SEED).integers(0, 13, (STEPS, N_TOTAL)).astype(np.int64)[:, base:base + n]
    run, hand = _run(dev, n, base, np.ascontiguousarray(acts), dist)
    if rank == 0:
        q.put((run, hand))
    dist.barrier()
    dist.destroy_process_group()


def test_two_ranks_reduce_to_one_engine():
    import torch
    import torch.multiprocessing as mp

    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29600 + os.getpid() % 200
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    run2, hand2 = q.get(timeout=600)
    for p in procs:
        p.join(timeout=600)
        assert p.exitcode == 0
    acts = np.random.default_rng(SEED).integers(0, 13, (STEPS, N_TOTAL)).astype(np.int64)
    run1, hand1 = _run(torch.device("cuda", 0), N_TOTAL, 0, acts)
    assert run2[0] == run1[0] == N_TOTAL * STEPS and run2[1] == run1[1]          # env steps, done flags
    np.testing.assert_allclose(run2[2:], run1[2:], rtol=1e-9)                        # float reward sums
    np.testing.assert_array_equal(hand2, hand1)                                      # integer hand statistics: exact
    assert hand1[0] > 0
