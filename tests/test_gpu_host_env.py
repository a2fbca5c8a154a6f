"""GPU suite: the host-buffer env (pulselib_b200/host_env.py, csrc/host_env.cu: byte actions in, compacted packed
records out, native threads expanding them) must hand back exactly what the float32-row engine returns -- observation
rows bit for bit, reward bits, done flags -- over whole hands, for every supported table shape, with batches that are
not multiples of the chunk / warp sizes, and against the oracle."""
import numpy as np
import pytest

from oracle import pyoracle

pytestmark = pytest.mark.gpu
CFG = dict(starting_bbs=100, max_bbs=1000, w1=.5, w2=.3, K=100, alpha=50)


@pytest.mark.parametrize("P,M,ap,N,chunks,threads", [(2, 2, 2, 50_000, 3, 4), (2, 2, 2, 1 << 20, 8, 0), (6, 6, 6, 33_333, 4, 3),
                                                      (6, 6, 3, 4097, 1, 2), (10, 10, 10, 20_001, 5, 0), (4, 4, 4, 999, 2, 1)])
def test_host_env_matches_float32_engine(P, M, ap, N, chunks, threads, cuda_device):
    import torch
    from unittest import mock

    from pulselib_b200.envs.poker_packed import PokerGPUPacked
    from pulselib_b200.host_env import PokerHostEnv

    seed, base = 20260401, 4242
    henv = PokerHostEnv(cuda_device, n_players=P, max_players=M, n_games=N, seed=seed, env_base=base, chunks=chunks,
                        threads=threads, **CFG)
    denv = PokerGPUPacked(cuda_device, n_players=P, max_players=M, n_games=N, seed=seed, env_base=base, **CFG)
    rng = np.random.default_rng(N)
    steps = 14 if P == 2 else 26
    for hand in range(3):
        rot = hand % P
        obs_h, _ = henv.reset(active_players=ap, rotation=rot)
        opts = {"rotation": rot}
        if ap != P:
            opts["active_players"] = True
            with mock.patch("torch.randint", lambda *a, **k: torch.tensor([ap])):
                obs_d, _ = denv.reset(options=opts)
        else:
            obs_d, _ = denv.reset(options=opts)
        np.testing.assert_array_equal(obs_h, obs_d.cpu().numpy(), err_msg=f"hand {hand} reset obs")
        assert not henv.dones.any() and not henv.rewards.any()
        for t in range(steps):
            a = rng.integers(0, 13, N)
            a_host = a.astype(np.uint8) if t % 2 else a.astype(np.int64)        # both action widths
            if t == 3:
                a_host = a.astype(np.int64); a_host[::7] = -1; a_host[1::7] = 13; a = a_host   # ids the reference ignores / treats alike
            o1, r1, d1, _, _ = henv.step(a_host)
            o2, r2, d2, _, _ = denv.step(torch.from_numpy(a.astype(np.int64)).to(cuda_device))
            tag = f"P{P} hand {hand} step {t}"
            np.testing.assert_array_equal(o1, o2.cpu().numpy(), err_msg=tag + " obs")
            np.testing.assert_array_equal(r1.view(np.uint32), r2.cpu().numpy().view(np.uint32), err_msg=tag + " reward bits")
            np.testing.assert_array_equal(d1, d2.cpu().numpy(), err_msg=tag + " done")
        inf = henv.info()
        assert inf["row_bytes"] == 4 * (3 + (13 + 3 * (M - 1) - 8 + 1) // 2)
        assert inf["last_d2h_bytes"] <= N * inf["row_bytes"] and inf["last_h2d_bytes"] >= N
    henv.close()


def test_host_env_matches_oracle(cuda_device, hr_table):
    from pulselib_b200.host_env import PokerHostEnv

    N = 200_000
    henv = PokerHostEnv(cuda_device, n_players=2, max_players=2, n_games=N, seed=5, env_base=99, chunks=4, **CFG)
    orc = pyoracle.OraclePoker(hr_table, n_players=2, max_players=2, n_games=N, **CFG)
    rng = np.random.default_rng(1)
    for hand in range(2):
        orc.tape_decks(5, 99, hand)
        np.testing.assert_array_equal(henv.reset()[0], orc.reset(None))
        for t in range(16):
            a = rng.integers(0, 13, N)
            o1, r1, d1, _, _ = henv.step(a.astype(np.uint8))
            o2, r2, d2 = orc.step(a)
            np.testing.assert_array_equal(o1, o2, err_msg=f"hand {hand} step {t}")
            np.testing.assert_array_equal(d1, d2)
            np.testing.assert_allclose(r1, r2, rtol=2e-6, atol=50 * 4e-7)
    henv.close()


def test_host_env_refuses_unrepresentable_configs(cuda_device):
    from pulselib_b200 import _lib
    from pulselib_b200.host_env import PokerHostEnv

    with pytest.raises(_lib.PulseError):
        PokerHostEnv(cuda_device, n_players=10, max_players=10, n_games=64, max_bbs=7000)   # 70,000 chips do not fit 16 bits
    with pytest.raises(_lib.PulseError):
        PokerHostEnv(cuda_device, n_players=3, max_players=10, n_games=64)                  # shape outside the specialised set


@pytest.mark.parametrize("P,M,N,chunks", [(2, 2, 70_001, 3), (6, 6, 9_999, 2)])
def test_records_mode_matches_float32_engine(P, M, N, chunks, cuda_device):
    """pulse_poker_host_step_records: the packed records the env leaves in pinned memory, decoded row by row with the
    library's own expander and scattered by the live mask, must rebuild exactly what the float32 engine returns."""
    import torch

    from pulselib_b200.envs.poker_packed import PokerGPUPacked
    from pulselib_b200.host_env import PokerHostEnv, expand_rows

    henv = PokerHostEnv(cuda_device, n_players=P, max_players=M, n_games=N, seed=8, env_base=17, chunks=chunks, threads=3, **CFG)
    denv = PokerGPUPacked(cuda_device, n_players=P, max_players=M, n_games=N, seed=8, env_base=17, **CFG)
    rng = np.random.default_rng(N)
    for hand in range(2):
        obs = henv.reset()[0].copy()
        rew = np.zeros(N, np.float32)
        np.testing.assert_array_equal(obs, denv.reset()[0].cpu().numpy())
        for t in range(18):
            a = rng.integers(0, 13, N)
            dones = henv.step_records(a.astype(np.uint8)).copy()
            live = henv.live_mask()
            rew[:] = 0
            for first, cnt, rows in henv.records():
                idx = first + np.flatnonzero(live[first:first + cnt])
                assert len(idx) == len(rows)
                o, r, d = expand_rows(np.array(rows), M)
                obs[idx], rew[idx] = o, r
                np.testing.assert_array_equal(d.astype(bool), dones[idx])
            o2, r2, d2, _, _ = denv.step(torch.from_numpy(a).to(cuda_device))
            np.testing.assert_array_equal(obs, o2.cpu().numpy(), err_msg=f"hand {hand} step {t}")
            np.testing.assert_array_equal(rew.view(np.uint32), r2.cpu().numpy().view(np.uint32))
            np.testing.assert_array_equal(dones, d2.cpu().numpy())
    henv.close()


# ---- 2048 with host buffers -----------------------------------------------------------------------------------------
@pytest.mark.parametrize("N,side,chunks,threads,auto", [(100_003, 4, 3, 4, True), (1 << 20, 4, 8, 0, True), (33_333, 3, 2, 2, True),
                                                         (5_000, 4, 4, 3, False), (257, 2, 1, 1, True)])
def test_tfe_host_env_matches_device_env(N, side, chunks, threads, auto, cuda_device):
    """TFEHostEnv (host arrays, packed boards over PCIe, native threads rebuilding the tile values) must return exactly
    what TFEBatched returns from device buffers: observations, rewards, done flags and scores, step for step."""
    import numpy as np
    import torch

    from pulselib_b200.envs.tfe import TFEBatched
    from pulselib_b200.host_env import TFEHostEnv

    ref = TFEBatched(cuda_device, N, side=side, seed=21, env_base=1 << 35, auto_reset=auto)
    henv = TFEHostEnv(cuda_device, n_envs=N, side=side, seed=21, env_base=1 << 35, auto_reset=auto, chunks=chunks, threads=threads)
    inf = henv.info()
    assert inf["chunks"] <= max(1, chunks) and inf["threads"] >= 1
    o, _ = ref.reset()
    ho, _ = henv.reset()
    np.testing.assert_array_equal(ho, o.cpu().numpy())
    assert not henv.dones.any()
    rng = np.random.default_rng(N)
    steps = 60 if N <= 200_000 else 12
    for t in range(steps):
        a = rng.integers(0, 4, N)
        a = a.astype(np.uint8) if t % 2 == 0 else a.astype(np.int64)
        o, r, d, _, info = ref.step(torch.from_numpy(a).to(cuda_device))
        ho, hr, hd, _, _ = henv.step(a)
        np.testing.assert_array_equal(ho, o.cpu().numpy(), err_msg=f"step {t}: observations")
        np.testing.assert_array_equal(hr, r.cpu().numpy(), err_msg=f"step {t}: rewards")
        np.testing.assert_array_equal(hd, d.cpu().numpy(), err_msg=f"step {t}: done flags")
    np.testing.assert_array_equal(henv.scores(), ref.total_score.cpu().numpy())
    np.testing.assert_array_equal(henv.boards().view(np.int64), ref.boards.cpu().numpy())
    assert henv.info()["last_d2h_bytes"] == 13 * N and henv.info()["last_h2d_bytes"] == N
    # a second hand on the same handle
    o, _ = ref.reset()
    ho, _ = henv.reset()
    np.testing.assert_array_equal(ho, o.cpu().numpy())
    henv.close()


def test_tfe_host_env_rejects_bad_arguments(cuda_device):
    import numpy as np

    from pulselib_b200 import _lib
    from pulselib_b200.host_env import TFEHostEnv

    with pytest.raises(ValueError):
        TFEHostEnv(cuda_device, n_envs=64, side=5)
    with pytest.raises(_lib.PulseError):
        TFEHostEnv(cuda_device, n_envs=0)
    env = TFEHostEnv(cuda_device, n_envs=64)
    with pytest.raises(_lib.PulseError):
        env.step(np.zeros(64, np.uint8))          # before reset
    env.reset()
    with pytest.raises(ValueError):
        env.step(np.zeros(63, np.uint8))


def test_host_env_reads_pinned_byte_actions_in_place(cuda_device):
    """a page-locked uint8 action array is copied from where it lies (no staging): same results as a pageable one"""
    import numpy as np

    from pulselib_b200.host_env import PokerHostEnv

    N = 70_001
    kw = dict(n_players=2, max_players=2, n_games=N, seed=5, env_base=3, chunks=3, threads=4, starting_bbs=100, w1=.5, w2=.3, K=100, alpha=50)
    a, b = PokerHostEnv(cuda_device, **kw), PokerHostEnv(cuda_device, **kw)
    acts = np.random.default_rng(1).integers(0, 13, (12, N)).astype(np.uint8)
    pinned = PokerHostEnv.pinned_bytes((12, N))
    pinned[...] = acts
    a.reset(); b.reset()
    for t in range(12):
        oa, ra, da, _, _ = a.step(pinned[t])
        ob, rb, db, _, _ = b.step(acts[t].copy())
        np.testing.assert_array_equal(oa, ob, err_msg=f"step {t}")
        np.testing.assert_array_equal(ra, rb)
        np.testing.assert_array_equal(da, db)
    assert a.timing()["stage_s"] < 0.25 * max(b.timing()["stage_s"], 1e-9) or a.timing()["stage_s"] < 1e-4
    a.close(); b.close()
