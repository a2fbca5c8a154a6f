"""GPU suite: the reference's own training loop and benchmark cases, replayed on the B200 through
`gym.make("Pulse-Poker-GPU-v1").unwrapped` -- the registered drop-in on the packed throughput layout.

tests/golden/poker_loops.npz was recorded in the build container from the live reference
(tests/golden/record_reference_loops.py): every env-facing call of scripts/Poker/trainGPU.py::train_agent (:37-110,
with environments/Poker/utils.py::build_actions and the reference's scripted agents) and of
benchmarking/Poker/cases.py run_env_reset / run_env_calculate_equities / run_env_execute_actions / run_env_step."""
import pytest
import torch

from tests import loops_util as lu

pytestmark = pytest.mark.gpu


def make_dropin(z, prefix, seed, device):
    from pulselib_b200 import registration
    from pulselib_b200.gym_compat import gym

    registration.register_all()
    return gym.make("Pulse-Poker-GPU-v1", device=device, agents=[], seed=seed, **lu.env_kwargs(z, prefix))


def test_train_loop_trace_calls(cuda_device):
    z = lu.load()
    env = make_dropin(z, "train", int(z["train/meta"][4]), cuda_device)
    assert lu.replay_calls(env, lu.calls_of(z, "train"), cuda_device, int(z["train/cfg"][3]), "train") >= 40
    assert env.unwrapped._mode == "packed"              # the loop never left the throughput layout
    env.close()


def test_train_loop_body(cuda_device):
    z = lu.load()
    env = make_dropin(z, "train", int(z["train/meta"][4]), cuda_device)
    lu.run_train_loop_body(env, z, cuda_device)
    assert env.unwrapped._mode == "packed"
    env.close()


def test_benchmark_cases(cuda_device):
    z = lu.load()
    lu.run_bench_cases(lambda seed: make_dropin(z, "bench", seed, cuda_device), z, cuda_device)


def test_info_seat_idx_compare_is_a_bool_tensor(cuda_device):
    """build_actions does `mask |= curr_players == seat_idx` (utils.py:113-115)."""
    from pulselib_b200.envs.poker import PokerGPU
    from pulselib_b200.envs.poker_packed import PokerGPUPacked

    for cls, kw in ((PokerGPU, dict(agents=[])), (PokerGPUPacked, {})):
        env = cls(cuda_device, n_players=3, max_players=3, n_games=4096, seed=3, **kw)
        obs, info = env.reset()
        m = info["seat_idx"] == 1
        assert torch.is_tensor(m) and m.dtype == torch.bool and tuple(m.shape) == (4096,) and m.device.type == "cuda"
        mask = torch.zeros_like(info["seat_idx"], dtype=torch.bool)
        mask |= info["seat_idx"] == 0
        assert bool(mask.any())
        assert tuple(info["stacks"][:, 1].shape) == (4096,)
