"""CPU, world_size 2 over gloo: the env-range split and the statistics all-reduce (the only collective of
the path), plus the property that makes sharding safe -- the chance tape depends on the global env id only."""
import os
import socket
import sys
from pathlib import Path

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

ROOT = Path(__file__).resolve().parents[1]


def test_shard_range_partitions_exactly():
    from pulselib_b200.sharding import shard_range

    for n, w in [(4_194_304, 8), (10, 3), (7, 8), (0, 2), (1_000_001, 4)]:
        parts = [shard_range(n, r, w) for r in range(w)]
        assert parts[0][0] == 0 and sum(c for _, c in parts) == n
        for (s0, c0), (s1, _) in zip(parts, parts[1:]):
            assert s0 + c0 == s1
        assert max(c for _, c in parts) - min(c for _, c in parts) <= 1
    with pytest.raises(ValueError):
        shard_range(10, 2, 2)


def _worker(rank, world, port, q):
    sys.path.insert(0, str(ROOT))
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port))
    from oracle import pyoracle
    from pulselib_b200 import sharding

    shard = sharding.init_from_env(n_total=1001, backend="gloo")
    # each rank steps its own slice of a 1001-board 2048 batch with the CPU oracle, chance from the tape
    orc = pyoracle.OracleTFE(shard.n_envs, 4)
    orc.reset(orc.tape(7, shard.env_base, pyoracle.STREAM_TFE_RESET, 0, 4))
    acts = np.random.default_rng(0).integers(0, 4, (20, 1001))
    stats = sharding.RunStats("cpu")
    for t in range(1, 21):
        a = acts[t - 1, shard.env_base: shard.env_base + shard.n_envs]
        _, rew, done = orc.step(a, orc.tape(7, shard.env_base, pyoracle.STREAM_TFE_SPAWN, t, 2))
        stats.vec += torch.tensor([shard.n_envs, float(done.sum()), float(rew.sum()), float((rew.astype(np.float64) ** 2).sum())],
                                  dtype=torch.float64)
    red = stats.all_reduce()
    q.put((rank, shard.env_base, shard.n_envs, red.tolist(), orc.boards.copy()))
    torch.distributed.destroy_process_group()


def test_two_rank_gloo_sharding_matches_single_process():
    from oracle import pyoracle

    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = sorted(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    # single-process run of all 1001 boards
    orc = pyoracle.OracleTFE(1001, 4)
    orc.reset(orc.tape(7, 0, pyoracle.STREAM_TFE_RESET, 0, 4))
    acts = np.random.default_rng(0).integers(0, 4, (20, 1001))
    want = np.zeros(4)
    for t in range(1, 21):
        _, rew, done = orc.step(acts[t - 1], orc.tape(7, 0, pyoracle.STREAM_TFE_SPAWN, t, 2))
        want += [1001, done.sum(), rew.sum(), (rew.astype(np.float64) ** 2).sum()]
    (r0, b0, n0, red0, boards0), (r1, b1, n1, red1, boards1) = got
    assert (b0, n0, b1, n1) == (0, 501, 501, 500)
    np.testing.assert_array_equal(np.concatenate([boards0, boards1]), orc.boards)   # shards == whole
    np.testing.assert_allclose(red0, want)
    np.testing.assert_allclose(red1, want)                                          # every rank holds the sum
