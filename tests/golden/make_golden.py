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


def main():
    cfg = dict(starting_bbs=100, w1=.5, w2=.3, K=100, alpha=50)  # config/pokerGPU.yaml:26-33
    poker_case("poker_hu", 2, 2, 48, 3, 16, 20260401, False, 0, cfg)
    poker_case("poker_6max", 6, 6, 32, 3, 24, 20260402, False, 1, cfg)
    poker_case("poker_10max_varap", 10, 10, 24, 4, 28, 20260403, True, 3, dict(cfg, starting_bbs=20, K=20, alpha=300, w2=.5))


if __name__ == "__main__":
    main()
