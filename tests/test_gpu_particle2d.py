"""GPU parity: Particle2D CUDA step vs the oracle restatement (rtol 1e-5 per BASELINE north_star; the
kernel and the oracle round every product before the add, so they normally agree bit for bit)."""
import numpy as np
import pytest

from oracle import pyoracle

pytestmark = pytest.mark.gpu


def test_particle2d_matches_oracle(cuda_device, oracle_lib):
    import ctypes as C
    import torch
    from pulselib_b200.envs.particle2d import Particle2D

    n = 100003
    env = Particle2D(cuda_device, n, dt=0.1, max_steps=50, seed=5)
    obs, _ = env.reset()
    state = obs.cpu().numpy().copy()
    assert np.all(state[:, 2:] == 0) and 4.5 < state[:, :2].std() < 5.5
    steps = np.zeros(n, np.int32)
    rng = np.random.default_rng(5)
    for t in range(60):
        act = rng.uniform(-1.5, 1.5, (n, 2)).astype(np.float32)
        obs, rew, term, trunc, _ = env.step(torch.from_numpy(act).to(cuda_device))
        r = np.zeros(n, np.float32)
        te = np.zeros(n, np.uint8)
        oracle_lib.oracle_particle2d_step(pyoracle.p(state), pyoracle.p(act), pyoracle.p(steps), pyoracle.p(r),
                                          pyoracle.p(te), C.c_float(0.1), C.c_int(50), C.c_int(n))
        np.testing.assert_allclose(obs.cpu().numpy(), state, rtol=1e-5, atol=1e-6)
        np.testing.assert_allclose(rew.cpu().numpy(), r, rtol=1e-5, atol=1e-6)
        np.testing.assert_array_equal(term.cpu().numpy(), te.astype(bool))
        np.testing.assert_array_equal(env.steps.cpu().numpy(), steps)
        assert not trunc.any()


def test_particle2d_rejects_cpu_device():
    from pulselib_b200._lib import PulseError
    from pulselib_b200.envs.particle2d import Particle2D

    with pytest.raises(PulseError):
        Particle2D("cpu", 4)
