"""Generates tests/golden/*.npz by running the reference's OWN env files from /root/reference on CPU
(build container only; the GPU box has no /root/reference).  Inputs (decks, actions, chance) and
every output the reference produced are stored, so the oracle and the CUDA kernels can be replayed
against them anywhere.

    python tests/golden/make_golden.py
"""
import sys
from pathlib import Path
from unittest import mock

import numpy as np
import torch

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
from oracle import ref_loader  # noqa: E402
from pulselib_b200 import handranks  # noqa: E402

OUT = Path(__file__).resolve().parent
INT_N = ("deck_positions", "pots", "stages", "button", "idx", "highest", "agg", "acted", "last_raise_size")
INT_NP = ("stacks", "current_round_bet", "total_invested", "status")


def poker_case(name, P, M, N, hands, steps, seed, vary_ap, rot, cfg):
    hr = handranks.load_table()
    rng = np.random.default_rng(seed)
    env = ref_loader.make_poker(hr, n_players=P, max_players=M, n_games=N, **cfg)
    rec = dict(meta=np.array([P, M, N, hands, steps], np.int32),
               cfg=np.array([cfg["starting_bbs"], cfg.get("max_bbs", 1000), cfg["K"], cfg["alpha"]], np.int32),
               w=np.array([cfg["w1"], cfg["w2"]], np.float32))
    decks_all, acts_all, ap_all, rot_all = [], [], [], []
    snaps = {k: [] for k in INT_N + INT_NP + ("board", "hands", "is_done", "obs", "rewards", "equities_pad")}
    reset_obs = []
    for h in range(hands):
        decks = (np.argsort(rng.random((N, 52)), axis=1) + 1).astype(np.int32)
        ap = int(rng.integers(2, P + 1)) if vary_ap else P
        r = (rot * h) % P
        opts = {"prefixed_decks": torch.from_numpy(decks), "rotation": r}
        if ap != P:
            opts["active_players"] = True
            with mock.patch("torch.randint", lambda *a, **k: torch.tensor([ap])):
                obs, _ = env.reset(options=opts)
        else:
            obs, _ = env.reset(options=opts)
        decks_all.append(decks); ap_all.append(ap); rot_all.append(r); reset_obs.append(obs.numpy().copy())
        acts = rng.integers(0, 13, (steps, N)).astype(np.int64)
        acts_all.append(acts)
        for t in range(steps):
            obs, rew, done, _, _ = env.step(torch.from_numpy(acts[t]))
            for k in INT_N + INT_NP + ("board", "hands"):
                snaps[k].append(getattr(env, k).numpy().astype(np.int32).copy())
            snaps["is_done"].append(done.numpy().copy())
            snaps["obs"].append(obs.numpy().copy())
            snaps["rewards"].append(rew.numpy().copy())
            eq = np.zeros((N, P), np.float32)
            eq[:, :ap] = env.equities.numpy()
            snaps["equities_pad"].append(eq)
    rec.update(decks=np.stack(decks_all), actions=np.stack(acts_all), active_players=np.array(ap_all, np.int32),
               rotation=np.array(rot_all, np.int32), reset_obs=np.stack(reset_obs))
    for k, v in snaps.items():
        rec[k] = np.stack(v).reshape((hands, steps) + v[0].shape)
    np.savez_compressed(OUT / f"{name}.npz", **rec)
    print(name, {k: v.shape for k, v in rec.items() if k in ("obs", "stacks")})


def blackjack_case(name, N, episodes, steps, seed):
    """Reference BlackJack (blackjack.py) on CPU.  Decks are injected through the reference's own seam:
    torch.rand is patched during reset so that argsort(rand) equals the wanted permutation (:28)."""
    from oracle import pyoracle

    mod = ref_loader.blackjack_module()
    env = mod.BlackJack(device=torch.device("cpu"), batch_size=N)
    orc = pyoracle.OracleBlackjack(N)
    rng = np.random.default_rng(seed)
    fields = ("deck_positions", "terminated", "players_cards", "players_card_idx", "player_card_sums", "has_ace",
              "dealer_cards", "dealer_card_idx", "dealer_upcard", "dealer_card_sums", "dealer_has_ace", "rewards", "obs")
    rec = {k: [] for k in fields}
    decks_all, acts_all, reset_obs = [], [], []
    for e in range(episodes):
        perm = orc.tape_decks(seed, 1000, e).copy()              # [N,52] permutation of 0..51 from the chance tape
        keys = np.empty((N, 52), np.float32)
        np.put_along_axis(keys, perm.astype(np.int64), np.arange(52, dtype=np.float32)[None, :] / 52.0, axis=1)
        with mock.patch("torch.rand", lambda *a, **k: torch.from_numpy(keys)):
            obs, _ = env.reset()
        assert np.array_equal(env.decks.numpy(), perm)
        decks_all.append(perm); reset_obs.append(obs.numpy().copy())
        acts = rng.integers(0, 2, (steps, N)).astype(np.int64)
        acts[:, : N // 8] = 0                                    # some envs hit until they bust
        acts_all.append(acts)
        for t in range(steps):
            env.step(torch.from_numpy(acts[t]))
            for k in fields:
                rec[k].append(getattr(env, k).numpy().copy())
    out = dict(meta=np.array([N, episodes, steps, seed, 1000], np.int64), decks=np.stack(decks_all),
               actions=np.stack(acts_all), reset_obs=np.stack(reset_obs))
    for k, v in rec.items():
        out[k] = np.stack(v).reshape((episodes, steps) + v[0].shape)
    np.savez_compressed(OUT / f"{name}.npz", **out)
    print(name, out["obs"].shape, "terminated at end:", out["terminated"][:, -1].mean())


def tfe_case(name, side, N, steps, seed):
    """Reference TFE (environments/2048/TFE.py): N independent single-board envs stepped in a Python loop.
    Chance is injected by replacing add_tile_numba in the loaded module with the tape rule of
    oracle/oracle_tfe.c (idx = mulhi32(u0, n_empty), four = u1 > floor(0.9 * 2**32))."""
    from oracle import pyoracle

    mod = ref_loader.tfe_module()
    orc = pyoracle.OracleTFE(N, side)
    draws = []

    def add_tile(board):
        u0, u1 = draws.pop(0)
        empties = [(r, c) for r in range(board.shape[0]) for c in range(board.shape[1]) if board[r, c] == 0]
        if not empties:
            return
        r, c = empties[(int(u0) * len(empties)) >> 32]
        board[r, c] = 4 if int(u1) > 3865470566 else 2

    rng = np.random.default_rng(seed)
    acts = rng.integers(0, 4, (steps, N)).astype(np.int64)
    reset_chance = orc.tape(seed, 500, pyoracle.STREAM_TFE_RESET, 0, 4)
    step_chance = np.stack([orc.tape(seed, 500, pyoracle.STREAM_TFE_SPAWN, t, 2) for t in range(steps)])
    boards = np.zeros((steps, N, side, side), np.int32)
    rewards = np.zeros((steps, N), np.int32)
    done = np.zeros((steps, N), np.uint8)
    scores = np.zeros((steps, N), np.int64)
    reset_boards = np.zeros((N, side, side), np.int32)
    with mock.patch.object(mod, "add_tile_numba", add_tile):
        for g in range(N):
            env = mod.TFE(side, side)
            draws[:] = [tuple(reset_chance[g, :2]), tuple(reset_chance[g, 2:])]
            obs, _ = env.reset()
            reset_boards[g] = obs
            for t in range(steps):
                draws[:] = [tuple(step_chance[t, g])]
                obs, rew, d, _, info = env.step(int(acts[t, g]))
                boards[t, g] = obs; rewards[t, g] = rew; done[t, g] = d; scores[t, g] = info["score"]
    np.savez_compressed(OUT / f"{name}.npz", meta=np.array([side, N, steps, seed, 500], np.int64), actions=acts,
                        reset_chance=reset_chance, step_chance=step_chance, reset_boards=reset_boards, boards=boards,
                        rewards=rewards, done=done, scores=scores)
    print(name, boards.shape, "done at end:", done[-1].mean(), "max tile:", boards.max())


def particle2d_case(name, N, steps, seed):
    """Reference Particle2D (Particle2D.py:15-30) on CPU with a given initial state and action tape."""
    mod = ref_loader.particle2d_module()
    env = mod.Particle2D(device=torch.device("cpu"), batch_size=N, dt=0.1, max_steps=40)
    rng = np.random.default_rng(seed)
    env.reset()
    init = np.zeros((N, 4), np.float32)
    init[:, :2] = (rng.standard_normal((N, 2)) * 5).astype(np.float32)
    init[: N // 4, :2] *= 0.01                                   # some particles start near the goal (dist < 0.1)
    env.state = torch.from_numpy(init.copy())
    acts = rng.uniform(-1.5, 1.5, (steps, N, 2)).astype(np.float32)  # beyond [-1,1]: the clamp matters
    st, rw, tm, sp = [], [], [], []
    for t in range(steps):
        s, r, term, trunc, _ = env.step(torch.from_numpy(acts[t]))
        st.append(s.numpy().copy()); rw.append(r.numpy().copy()); tm.append(term.numpy().copy()); sp.append(env.steps.numpy().copy())
    np.savez_compressed(OUT / f"{name}.npz", meta=np.array([N, steps, 40], np.int64), dt=np.float32(0.1), init=init,
                        actions=acts, state=np.stack(st), reward=np.stack(rw), terminated=np.stack(tm), steps=np.stack(sp))
    print(name, np.stack(st).shape, "terminated frac:", np.stack(tm).mean())


def main():
    cfg = dict(starting_bbs=100, w1=.5, w2=.3, K=100, alpha=50)  # config/pokerGPU.yaml:26-33
    cases = {
        "poker_hu": lambda: poker_case("poker_hu", 2, 2, 48, 3, 16, 20260401, False, 0, cfg),
        "poker_6max": lambda: poker_case("poker_6max", 6, 6, 32, 3, 24, 20260402, False, 1, cfg),
        "poker_10max_varap": lambda: poker_case("poker_10max_varap", 10, 10, 24, 4, 28, 20260403, True, 3,
                                                dict(cfg, starting_bbs=20, K=20, alpha=300, w2=.5)),
        "blackjack": lambda: blackjack_case("blackjack", 512, 3, 12, 1),
        "tfe_4x4": lambda: tfe_case("tfe_4x4", 4, 48, 400, 2),
        "tfe_3x3": lambda: tfe_case("tfe_3x3", 3, 32, 120, 3),
        "particle2d": lambda: particle2d_case("particle2d", 1024, 60, 5),
    }
    for name in (sys.argv[1:] or list(cases)):   # python tests/golden/make_golden.py [case ...]
        cases[name]()


if __name__ == "__main__":
    main()
