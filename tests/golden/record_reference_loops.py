"""Records what the reference's OWN training loop and benchmark cases do to the reference env (build container only).

`scripts/Poker/trainGPU.py::train_agent` (:37-110) and `benchmarking/Poker/cases.py::run_env_*` (:66-170) are imported
IN PLACE from /root/reference and run on CPU against the reference's `environments/Poker/PokerGPU.py`, with the
reference's own `build_actions` (environments/Poker/utils.py:108-123) and scripted agents
(environments/Poker/Player.py:79-176).  A recording proxy sits where `gym.make(...)` would hand out the env and
stores every env-facing call -- reset options, step actions, and everything that came back (obs, rewards, dones,
info["seat_idx"], info["stacks"], info["active_players"]) -- into tests/golden/poker_loops.npz, together with what the
loop itself computed from them (Q-seat transition counts per step, episode reward, episode profit, step count).

The replays (tests/test_reference_loops.py on the oracle, tests/test_gpu_reference_loops.py on the B200 through
`gym.make("Pulse-Poker-GPU-v1")`) make the same calls and must get the same answers, then run the loop body itself.

Chance: the trainer's reset passes no decks (the reference draws `rand().argsort()`), so the proxy injects the chance
tape's decks -- tape(seed, table, hand) -- through the reference's own `prefixed_decks` seam; the engine draws the very
same decks natively.  The table-size draw `torch.randint(2, P + 1)` (:76-80) is recorded and replayed.  The Q agent is
a stand-in that plays a seeded action tape (a network would not be bit-reproducible across devices); the scripted
opponents draw from torch's seeded CPU generator -- only the actions they produced are kept.

    python tests/golden/record_reference_loops.py
"""
from __future__ import annotations

import sys
import tempfile
import types
from pathlib import Path
from unittest import mock

import numpy as np
import torch

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
REF = Path("/root/reference")
OUT = Path(__file__).resolve().parent / "poker_loops.npz"
SEED = 20260405


def install_standins():
    from pulselib_b200 import gym_compat

    gym_compat.install_shim()
    ev = types.ModuleType("eval7")
    ev.Card = lambda s: None
    ev.evaluate = lambda cards: 0
    ev.HandRange = lambda s: None
    ev.Deck = lambda: None
    sys.modules["eval7"] = ev
    mpl = types.ModuleType("matplotlib")
    mpl.use = lambda *a, **k: None
    plt = types.ModuleType("matplotlib.pyplot")
    mpl.pyplot = plt
    sys.modules["matplotlib"], sys.modules["matplotlib.pyplot"] = mpl, plt
    # bare packages so the reference's modules load by file without its package __init__s
    # (environments/Poker/__init__.py pulls in Poker.py, which needs the real eval7)
    for name, rel in (("environments", "environments"), ("environments.Poker", "environments/Poker")):
        m = types.ModuleType(name)
        m.__path__ = [str(REF / rel)]
        sys.modules[name] = m
    sys.path.insert(1, str(REF))


def make_reference_env(pokergpu_mod, hand_ranks_np, **kw):
    table = torch.from_numpy(hand_ranks_np)

    class _Stat:
        st_size = table.numel() * 4

    with mock.patch.object(Path, "exists", lambda self: True), \
            mock.patch.object(Path, "stat", lambda self, **k: _Stat()), \
            mock.patch.object(torch, "from_file", lambda **k: table):
        return pokergpu_mod.PokerGPU(device=torch.device("cpu"), **kw)


def npy(t):
    a = t.detach().cpu().numpy() if torch.is_tensor(t) else np.asarray(t)
    return np.array(a.astype(np.uint8) if a.dtype == np.bool_ else a, copy=True)


class RecordingEnv:
    """Stands where gym.make() would hand out the env: forwards everything to the reference env, records reset / step."""

    def __init__(self, env, oracle_for_decks, seed, inject_decks=True):
        self.__dict__.update(_env=env, _orc=oracle_for_decks, _seed=seed, _hand=0, calls=[], _inject=inject_decks)

    @property
    def unwrapped(self):
        return self._env

    def __getattr__(self, name):
        return getattr(self._env, name)

    def close(self):
        return None

    def reset(self, **kw):
        options = dict(kw.get("options") or {})
        rec = dict(kind="reset", rotation=int(options.get("rotation", 0)), q_agent_seat=int(options.get("q_agent_seat", 0)),
                   active_players_flag=int(bool(options.get("active_players", False))), hand=self._hand)
        if self._inject:
            self._orc.tape_decks(self._seed, 0, self._hand)
            options["prefixed_decks"] = torch.from_numpy(self._orc.decks.copy())
        drawn = []
        real_randint = torch.randint

        def spy(*a, **k):
            out = real_randint(*a, **k)
            drawn.append(int(out.reshape(-1)[0]))
            return out

        with mock.patch("torch.randint", spy):
            obs, info = self._env.reset(options=options)
        rec["ap_draw"] = drawn[0] if drawn else -1
        rec.update(obs=npy(obs), seat_idx=npy(info["seat_idx"]), stacks=npy(info["stacks"]),
                   active_players=int(info["active_players"]), button=npy(self._env.button), decks=npy(self._env.decks))
        self._hand += 1
        self.calls.append(rec)
        return obs, info

    def step(self, actions):
        a = npy(actions)
        obs, rew, done, trunc, info = self._env.step(actions)
        self.calls.append(dict(kind="step", actions=a, obs=npy(obs), rewards=npy(rew), dones=npy(done),
                               seat_idx=npy(info["seat_idx"]), stacks=npy(info["stacks"]),
                               active_players=int(info["active_players"])))
        return obs, rew, done, trunc, info


class TapeQAgent:
    """Stand-in for PokerQNetwork: plays a seeded action tape, records the transition batches it is handed."""

    def __init__(self, seed):
        self.rng = np.random.default_rng(seed)
        self.batches = []
        self.network = types.SimpleNamespace(state_dict=lambda: {})

    def get_actions(self, states):
        return torch.from_numpy(self.rng.integers(0, 13, states.shape[0]).astype(np.int64))

    def train_step(self, states, actions, rewards, next_states, dones):
        self.batches.append((int(states.shape[0]), float(rewards.double().sum()), int(dones.sum()),
                             float(states.double().sum()), float(next_states.double().sum()), int(actions.sum())))


class Plots:
    def __init__(self):
        self.curves = []

    def plot_learning_curve(self, *, scores, file_path, window_size, title):
        self.curves.append([float(s) for s in scores])


class NoBench:
    def create_benchmark_file(self, **kw):
        self.kw = kw


def pack_calls(prefix, calls, out):
    out[f"{prefix}/n_calls"] = np.int32(len(calls))
    out[f"{prefix}/kinds"] = np.array([c["kind"] for c in calls])
    for i, c in enumerate(calls):
        for k, v in c.items():
            if k != "kind":
                out[f"{prefix}/c{i}/{k}"] = np.asarray(v)


def record_train_loop(out):
    from oracle import pyoracle
    from pulselib_b200 import handranks
    import environments.Poker.PokerGPU as pokergpu
    import environments.Poker.utils as putils
    from scripts.Poker import trainGPU as train_gpu

    hr = handranks.load_table()
    N, names = 512, ["tight_aggressive", "heuristic_hands", "loose_passive"]
    cfg = dict(starting_bbs=100, w1=.5, w2=.3, K=100, alpha=50)          # config/pokerGPU.yaml:26-33
    dev = torch.device("cpu")
    torch.manual_seed(SEED)
    agents, agent_types = putils.load_gpu_agents(dev, len(names), names, cfg["starting_bbs"], 13)
    q = TapeQAgent(SEED)
    agents.insert(0, q)
    agent_types.insert(0, putils.PokerAgentType.QLEARNING)                 # scripts/Poker/trainGPU.py:174-175
    P = len(names) + 1
    env = make_reference_env(pokergpu, hr, agents=agents, n_players=P, n_games=N, **cfg)   # :177-188 (max_players 10)
    orc = pyoracle.OraclePoker(hr, n_players=P, max_players=10, n_games=N, **cfg)
    rec_env = RecordingEnv(env, orc, SEED)
    plots, episodes = Plots(), 4
    with tempfile.TemporaryDirectory() as tmp:
        train_gpu.train_agent(rec_env, agents, agent_types, episodes, N, device=dev, results_dir=Path(tmp),
                              config={"RESULTS_DIR": "x"}, plotter=plots, benchmarker=NoBench())
    pack_calls("train", rec_env.calls, out)
    out["train/meta"] = np.array([P, 10, N, episodes, SEED], np.int64)
    out["train/cfg"] = np.array([cfg["starting_bbs"], 1000, cfg["K"], cfg["alpha"]], np.int32)
    out["train/w"] = np.array([cfg["w1"], cfg["w2"]], np.float32)
    out["train/agent_kinds"] = np.array(["qlearning"] + names)
    out["train/reward_scores"] = np.array(plots.curves[0], np.float64)
    out["train/profit_scores"] = np.array(plots.curves[1], np.float64)
    out["train/q_batches"] = np.array(q.batches, np.float64)
    steps = sum(1 for c in rec_env.calls if c["kind"] == "step")
    print(f"train loop: {episodes} episodes, {steps} env steps recorded, profits {plots.curves[1]}")


def record_benchmark_cases(out):
    """benchmarking/Poker/cases.py run_env_reset / run_env_calculate_equities / run_env_execute_actions / run_env_step
    with warmup 1, measure 2, N_GAMES 256, on the reference env; the white-box cases also store the state they left."""
    from oracle import pyoracle
    from pulselib_b200 import handranks
    import yaml
    import environments.Poker.PokerGPU as pokergpu
    import environments.Poker.utils as putils
    from benchmarking.Poker import cases, runtime

    hr = handranks.load_table()
    config = yaml.safe_load((REF / "config" / "pokerGPU.yaml").read_text())
    bcfg = dict(config, N_GAMES=256, EPISODES=1)
    dev = torch.device("cpu")
    ctx = runtime.BenchmarkContext(config=config, benchmark_config=bcfg, device=dev, root_dir=ROOT,
                                   results_dir=Path(tempfile.gettempdir()), weights_path=Path(tempfile.gettempdir()) / "w.pth")
    P = bcfg["NUM_PLAYERS"] + 1
    envs = []

    def create_agents_and_types(context):
        agents, types_ = putils.load_gpu_agents(dev, bcfg["NUM_PLAYERS"], bcfg["AGENTS"], bcfg["STARTING_BBS"], 13)
        q = TapeQAgent(SEED + 1)
        agents.insert(0, q)
        types_.insert(0, putils.PokerAgentType.QLEARNING)
        return agents, types_, q

    def create_env(context, agents):                                     # runtime.create_env (:104-117) on CPU
        env = make_reference_env(pokergpu, hr, agents=agents, n_players=P, n_games=bcfg["N_GAMES"],
                                 starting_bbs=bcfg["STARTING_BBS"], w1=bcfg["W1"], w2=bcfg["W2"], K=bcfg["K"], alpha=bcfg["ALPHA"])
        orc = pyoracle.OraclePoker(hr, n_players=P, max_players=10, n_games=bcfg["N_GAMES"])
        rec = RecordingEnv(env, orc, SEED + 7 + len(envs))
        envs.append(rec)
        return rec

    STATE = ("deck_positions", "pots", "stages", "button", "sb", "bb", "idx", "highest", "agg", "acted", "last_raise_size",
             "stacks", "current_round_bet", "total_invested", "status", "board", "hands", "equities", "equity_dirty", "decks")

    def snap(env):
        return {k: npy(getattr(env, k)) for k in STATE}

    case = types.SimpleNamespace(name="x", category="env", description="", primary_metric_name="latency",
                                 primary_metric_unit="seconds", lower_is_better=True)
    torch.manual_seed(SEED + 1)
    with mock.patch.object(runtime, "create_agents_and_types", create_agents_and_types), \
            mock.patch.object(runtime, "create_env", create_env):
        for name in ("run_env_reset", "run_env_calculate_equities", "run_env_execute_actions", "run_env_step"):
            before = len(envs)
            res = getattr(cases, name)(case, ctx, 1, 2)
            rec = envs[before]
            pack_calls(f"bench/{name}", rec.calls, out)
            out[f"bench/{name}/seed"] = np.int64(rec._seed)
            for k, v in snap(rec.unwrapped).items():
                out[f"bench/{name}/final/{k}"] = v
            out[f"bench/{name}/final/active_players"] = np.int32(rec.unwrapped.active_players)
            print(f"{name}: {len(rec.calls)} env calls, mean {res['primary_metric']['value'] * 1e3:.2f} ms on the reference (CPU)")
    out["bench/meta"] = np.array([P, 10, bcfg["N_GAMES"]], np.int64)
    out["bench/cfg"] = np.array([bcfg["STARTING_BBS"], 1000, bcfg["K"], bcfg["ALPHA"]], np.int32)
    out["bench/w"] = np.array([bcfg["W1"], bcfg["W2"]], np.float32)


def main():
    install_standins()
    out = {}
    record_train_loop(out)
    record_benchmark_cases(out)
    np.savez_compressed(OUT, **out)
    print(OUT, f"{OUT.stat().st_size / 1e6:.2f} MB")


if __name__ == "__main__":
    main()
