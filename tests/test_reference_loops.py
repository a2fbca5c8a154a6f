"""CPU suite: the drop-in class `Pulse-Poker-GPU-v1` (pulselib_b200/envs/poker.py::PokerGPU -- packed engine, mirror on
access) driven by the traces recorded from the reference's own training loop and benchmark cases
(tests/golden/record_reference_loops.py).  No GPU here: the class runs on the host-compiled kernel bodies through
tests/hostemu/shim.py, which exercises exactly the host logic the GPU run uses (mode switching, version-counter write
detection, live info tensors); tests/test_gpu_reference_loops.py makes the same replays on the B200."""
import numpy as np
import pytest
import torch

from tests import loops_util as lu
from tests.hostemu.shim import hostemu_backend


def make_dropin(z, prefix, seed):
    from pulselib_b200 import registration
    from pulselib_b200.gym_compat import gym

    registration.register_all()
    return gym.make("Pulse-Poker-GPU-v1", device="cpu", agents=[], seed=seed, **lu.env_kwargs(z, prefix))


def test_train_loop_trace_calls():
    z = lu.load()
    with hostemu_backend():
        env = make_dropin(z, "train", int(z["train/meta"][4]))
        n = lu.replay_calls(env, lu.calls_of(z, "train"), torch.device("cpu"), int(z["train/cfg"][3]), "train")
        assert n >= 40
        assert env.unwrapped._mode == "packed"          # the loop never left the throughput layout
        env.close()


def test_train_loop_body():
    z = lu.load()
    with hostemu_backend():
        env = make_dropin(z, "train", int(z["train/meta"][4]))
        lu.run_train_loop_body(env, z, torch.device("cpu"))
        assert env.unwrapped._mode == "packed"
        env.close()


def test_benchmark_cases():
    z = lu.load()
    with hostemu_backend():
        lu.run_bench_cases(lambda seed: make_dropin(z, "bench", seed), z, torch.device("cpu"))


def test_info_tensors_are_real_and_live():
    """`info["seat_idx"] == q_seat` must be a bool tensor, not the Python bool False."""
    with hostemu_backend():
        from pulselib_b200.envs.poker import PokerGPU
        from pulselib_b200.envs.poker_packed import PokerGPUPacked

        for cls, kw in ((PokerGPU, dict(agents=[])), (PokerGPUPacked, {})):
            env = cls("cpu", n_players=3, max_players=3, n_games=64, seed=3, **kw)
            obs, info = env.reset()
            m = info["seat_idx"] == 1
            assert torch.is_tensor(m) and m.dtype == torch.bool and tuple(m.shape) == (64,)
            assert torch.is_tensor(info["stacks"]) and tuple(info["stacks"].shape) == (64, 3)
            first = info["stacks"][:, 0].clone()
            seat0 = info["seat_idx"].clone()
            for _ in range(6):
                obs, rew, done, trunc, info2 = env.step(torch.full((64,), 8, dtype=torch.long))
            assert not torch.equal(info["stacks"][:, 0], first) or not torch.equal(info["seat_idx"], seat0)  # live, not snapshots
            np.testing.assert_array_equal(info["seat_idx"].numpy(), info2["seat_idx"].numpy())


def test_mirror_on_access_then_write_switches_layout():
    with hostemu_backend():
        from pulselib_b200.envs.poker import PokerGPU

        env = PokerGPU("cpu", agents=[], n_players=4, max_players=4, n_games=32, seed=11)
        ref = PokerGPU("cpu", agents=[], n_players=4, max_players=4, n_games=32, seed=11, layout="compat")
        env.reset(); ref.reset()
        a = torch.ones(32, dtype=torch.long)
        env.step(a); ref.step(a)
        assert env._mode == "packed"
        for f in ("stages", "board", "status", "pots", "decks", "equities", "equity_dirty", "sb", "bb", "hands",
                  "current_round_bet", "total_invested", "deck_positions", "prev_invested", "button", "is_done"):
            np.testing.assert_array_equal(getattr(env, f).numpy(), getattr(ref, f).numpy(), err_msg=f)
        assert env._mode == "packed"                      # reads alone change nothing
        env.step(a); ref.step(a)
        assert env._mode == "packed" and "stages" not in env.__dict__   # the snapshot was dropped by the step
        env.stacks[0] = torch.tensor([7, 9, 11, 13], dtype=torch.int32)  # a white-box write, as the reference's tests do
        ref.stacks[0] = torch.tensor([7, 9, 11, 13], dtype=torch.int32)
        env.pots[1] = 55; ref.pots[1] = 55
        o1, r1, d1, _, i1 = env.step(a)
        o2, r2, d2, _, i2 = ref.step(a)
        assert env._mode == "compat"
        np.testing.assert_array_equal(o1.numpy(), o2.numpy())
        np.testing.assert_array_equal(i1["stacks"].numpy(), i2["stacks"].numpy())
        env.reset(); ref.reset()                          # back on the throughput layout, stacks and button carried over
        assert env._mode == "packed"
        for t in range(5):
            o1, r1, d1, _, i1 = env.step(a * (t % 3))
            o2, r2, d2, _, i2 = ref.step(a * (t % 3))
            np.testing.assert_array_equal(o1.numpy(), o2.numpy())
            np.testing.assert_array_equal(i1["stacks"].numpy(), i2["stacks"].numpy())
            np.testing.assert_array_equal(env.button.numpy(), ref.button.numpy())
