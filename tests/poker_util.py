"""Shared replay/compare helpers for the poker parity tests."""
from pathlib import Path

import numpy as np

GOLDEN = Path(__file__).resolve().parent / "golden"
INT_N = ("deck_positions", "pots", "stages", "button", "idx", "highest", "agg", "acted", "last_raise_size")
INT_NP = ("stacks", "current_round_bet", "total_invested", "status")
# |alpha * tanh| differs between libm, CUDA and Sleef tanhf by a few ulp of tanh (<= ~2.4e-7 each)
REWARD_RTOL = 2e-6


def reward_atol(alpha):
    return abs(alpha) * 4e-7


def to_np(x):
    if hasattr(x, "detach"):
        return x.detach().cpu().numpy()
    return np.asarray(x)


class TorchAdapter:
    """Gives pulselib_b200.envs.poker.PokerGPU the (decks, active_players, rotation) reset signature."""

    def __init__(self, env):
        self.env = env

    def reset(self, decks, active_players=None, rotation=0):
        import torch
        from unittest import mock

        opts = {"rotation": rotation}
        if decks is not None:
            opts["prefixed_decks"] = torch.from_numpy(np.ascontiguousarray(decks))
        P = self.env.n_players
        if active_players is not None and active_players != P:
            opts["active_players"] = True
            with mock.patch("torch.randint", lambda *a, **k: torch.tensor([active_players])):
                obs, _ = self.env.reset(options=opts)
        else:
            obs, _ = self.env.reset(options=opts)
        return obs

    def step(self, actions):
        import torch

        obs, rew, done, _, _ = self.env.step(torch.from_numpy(np.ascontiguousarray(actions)).to(self.env.device))
        self.rewards = rew
        return obs, rew, done

    def __getattr__(self, name):
        return getattr(self.env, name)


def compare_state(a, b, tag, fields=INT_N + INT_NP + ("board", "hands")):
    for f in fields:
        x, y = to_np(getattr(a, f)).astype(np.int64), to_np(getattr(b, f)).astype(np.int64)
        if not np.array_equal(x, y):
            bad = tuple(np.argwhere(x != y)[0])
            raise AssertionError(f"{tag}: field {f} differs at {bad}: {x[bad]} vs {y[bad]}")


def replay_golden(name, engine, check_equities=True):
    """Replays a golden file through `engine`; asserts bit-exact ints/obs/equities, reward within tolerance."""
    z = np.load(GOLDEN / f"{name}.npz")
    P, M, N, hands, steps = z["meta"]
    alpha = int(z["cfg"][3])
    for h in range(hands):
        ap = int(z["active_players"][h])
        obs = engine.reset(z["decks"][h], active_players=ap, rotation=int(z["rotation"][h]))
        np.testing.assert_array_equal(to_np(obs), z["reset_obs"][h], err_msg=f"{name} reset obs hand {h}")
        for t in range(steps):
            obs, rew, done = engine.step(z["actions"][h, t])
            tag = f"{name} hand {h} step {t}"
            for f in INT_N + INT_NP + ("board", "hands"):
                got = to_np(getattr(engine, f)).astype(np.int32)
                if not np.array_equal(got, z[f][h, t]):
                    bad = tuple(np.argwhere(got != z[f][h, t])[0])
                    raise AssertionError(f"{tag}: {f} at {bad}: got {got[bad]} want {z[f][h, t][bad]}")
            np.testing.assert_array_equal(to_np(done).astype(bool), z["is_done"][h, t], err_msg=tag + " done")
            np.testing.assert_array_equal(to_np(obs), z["obs"][h, t], err_msg=tag + " obs")
            if check_equities:
                np.testing.assert_array_equal(to_np(engine.equities), z["equities_pad"][h, t][:, :ap], err_msg=tag + " equities")
            np.testing.assert_allclose(to_np(rew), z["rewards"][h, t], rtol=REWARD_RTOL, atol=reward_atol(alpha),
                                       err_msg=tag + " reward")


def golden_kwargs(name):
    z = np.load(GOLDEN / f"{name}.npz")
    P, M, N, _, _ = (int(v) for v in z["meta"])
    sb, mb, K, alpha = (int(v) for v in z["cfg"])
    return dict(n_players=P, max_players=M, n_games=N, starting_bbs=sb, max_bbs=mb, K=K, alpha=alpha,
                w1=float(z["w"][0]), w2=float(z["w"][1]))


# every reference attribute the packed layout's mirror rebuilds (pulse_poker_unpack_full)
MIRROR_FIELDS = INT_N + INT_NP + ("board", "hands", "sb", "bb", "prev_invested", "decks", "equity_dirty")


def random_rollout_compare(a, b, rng, P, N, hands, steps, vary_ap=False, rot=0, alpha=50, check_equities=True,
                           fields=INT_N + INT_NP + ("board", "hands")):
    """Runs engines a and b on the same decks/actions; all integer state must match exactly."""
    for h in range(hands):
        ap = int(rng.integers(2, P + 1)) if vary_ap else P
        decks = (np.argsort(rng.random((N, 52)), axis=1) + 1).astype(np.int32)
        r = (rot * h) % P
        oa = a.reset(decks, active_players=ap, rotation=r)
        ob = b.reset(decks, active_players=ap, rotation=r)
        np.testing.assert_array_equal(to_np(oa), to_np(ob), err_msg=f"reset obs hand {h}")
        compare_state(a, b, f"reset hand {h}", fields)
        for t in range(steps):
            acts = rng.integers(0, 13, N).astype(np.int64)
            oa, ra, da = a.step(acts)
            ob, rb, db = b.step(acts)
            tag = f"hand {h} step {t}"
            compare_state(a, b, tag, fields)
            np.testing.assert_array_equal(to_np(da).astype(bool), to_np(db).astype(bool), err_msg=tag + " done")
            np.testing.assert_array_equal(to_np(oa), to_np(ob), err_msg=tag + " obs")
            if check_equities:
                np.testing.assert_array_equal(to_np(a.equities), to_np(b.equities), err_msg=tag + " equities")
            np.testing.assert_allclose(to_np(ra), to_np(rb), rtol=REWARD_RTOL, atol=reward_atol(alpha), err_msg=tag + " reward")
            if to_np(da).all():
                break


# ---- replay of the call traces recorded from the reference's own unit tests -------------------------------
# tests/golden/poker_reftests.npz (made by tests/golden/record_reference_tests.py): every top-level env call
# the reference's tests/poker/test_poker_gpu_*.py make on the reference env, with full state before/after.
KAT_N = ("deck_positions", "pots", "stages", "button", "sb", "bb", "idx", "highest", "agg", "acted",
         "last_raise_size", "prev_invested")
KAT_NP = ("stacks", "current_round_bet", "total_invested", "status")
KAT_OTHER = ("decks", "board", "hands", "is_done", "equity_dirty", "equities")


class KatRecord:
    def __init__(self, z, i):
        self.i, self.test, self.method = i, str(z["tests"][i]), str(z["methods"][i])
        self.env_id, self.first, self.had_button = int(z["env_ids"][i]), bool(z["first_reset"][i]), bool(z["had_button"][i])
        pre = f"r{i}/"
        self.ctor, self.args, self.before, self.after, self.ret = {}, {}, {}, {}, {}
        for k in z.files:
            if k.startswith(pre):
                _, part, name = k.split("/")
                getattr(self, part)[name] = z[k]
        self.ctor = {k: (float(v) if k in ("w1", "w2") else int(v)) for k, v in self.ctor.items()}


def load_reference_kat():
    z = np.load(GOLDEN / "poker_reftests.npz")
    return [KatRecord(z, i) for i in range(int(z["n_records"]))]


def replay_reference_kat(make_engine, records=None):
    """make_engine(ctor_kwargs) -> adapter with set_state(dict, active_players), get_state(name),
    reset(rec) -> obs, call(rec) -> dict of returned arrays.  Every record is replayed on a fresh engine:
    set "before", make the call, compare with "after" (ints/obs exact, fp32 reward within tolerance)."""
    records = records if records is not None else load_reference_kat()
    checked = 0
    for r in records:
        eng = make_engine(r.ctor)
        tag = f"[{r.i}] {r.test} :: {r.method}"
        eng.set_state(r.before, int(r.before["active_players"]), r)
        ret = eng.reset(r) if r.method == "reset" else eng.call(r)
        assert int(eng.get_state("active_players")) == int(r.after["active_players"]), tag + " active_players"
        for f in KAT_N + KAT_NP + KAT_OTHER:
            if f not in r.after:
                continue
            want, got = r.after[f], np.asarray(eng.get_state(f))
            if f == "equities":
                ap = int(r.after["active_players"])   # columns >= active_players are never read (see set_state)
                np.testing.assert_array_equal(got.reshape(want.shape[0], -1)[:, :ap], want[:, :ap], err_msg=tag + " equities")
                continue
            got = got.astype(want.dtype).reshape(want.shape)
            if not np.array_equal(got, want):
                bad = tuple(np.argwhere(got != want)[0])
                raise AssertionError(f"{tag}: {f} at {bad}: got {got[bad]} want {want[bad]}")
        alpha = abs(r.ctor["alpha"])
        for k, want in r.ret.items():
            got = np.asarray(ret[k])
            if k == "reward" or (k == "value" and r.method == "poker_reward_gpu"):
                np.testing.assert_allclose(got, want, rtol=REWARD_RTOL, atol=reward_atol(alpha), err_msg=tag + " " + k)
            else:
                np.testing.assert_array_equal(got.astype(want.dtype).reshape(want.shape), want, err_msg=tag + " ret " + k)
        checked += 1
    return checked
