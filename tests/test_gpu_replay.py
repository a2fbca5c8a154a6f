"""GPU suite for the transition hand-off row: the device ring against the trace recorded from the
reference's own ReplayBufferTorch and against the numpy restatement on ragged sizes."""
import numpy as np
import pytest
import torch

from oracle import pyoracle
from tests.test_replay_cpu import Z, check_ring, replay_trace

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("tag", ["wrap", "six"])
def test_device_ring_matches_reference_buffer(tag, cuda_device):
    from pulselib_b200.replay import TransitionRing

    N, W, P, q_seat, steps, cap = (int(v) for v in Z[f"{tag}/meta"])
    ring = TransitionRing(cuda_device, cap, W)

    def append(state, actions, rewards, next_state, dones, seat, term_before, q):
        t = [torch.from_numpy(np.ascontiguousarray(a)).to(cuda_device) for a in (state, actions, rewards, next_state, dones, seat, term_before)]
        ring.append(t[0], t[1], t[2], t[3], t[4], t[5], t[6], q)

    replay_trace(tag, append)
    check_ring(tag, ring.state_buffer.cpu().numpy(), ring.next_state_buffer.cpu().numpy(), ring.action_buffer.cpu().numpy(),
               ring.reward_buffer.cpu().numpy(), ring.terminated_buffer.cpu().numpy(), ring.position, len(ring))


@pytest.mark.parametrize("N,W,cap", [(100_003, 16, 1 << 18), (70_001, 40, 70_001), (257, 13, 1000)])
def test_device_ring_matches_oracle_ragged(N, W, cap, cuda_device):
    from pulselib_b200.replay import TransitionRing

    rng = np.random.default_rng(N)
    ring, orc = TransitionRing(cuda_device, cap, W), pyoracle.OracleRing(cap, W)
    term = np.zeros(N, bool)
    for t in range(7):
        state = rng.standard_normal((N, W)).astype(np.float32)
        nxt = rng.standard_normal((N, W)).astype(np.float32)
        acts, rew = rng.integers(0, 13, N), rng.standard_normal(N).astype(np.float32)
        seat = rng.integers(0, 3, N).astype(np.int32)
        dones = (rng.random(N) < 0.15) | term
        use_term = None if t == 3 else term
        k = orc.append(state, acts, rew, nxt, dones, seat, use_term, 1)
        tt = [torch.from_numpy(a).to(cuda_device) for a in (state, acts, rew, nxt, dones, seat)]
        ring.append(tt[0], tt[1], tt[2], tt[3], tt[4], tt[5], None if use_term is None else torch.from_numpy(term).to(cuda_device), 1)
        assert int(ring.appended.item()) == k
        term = term | dones
    assert (ring.position, len(ring)) == (orc.position, orc.size)
    for name in ("state_buffer", "next_state_buffer", "action_buffer", "reward_buffer", "terminated_buffer"):
        np.testing.assert_array_equal(getattr(ring, name).cpu().numpy(), getattr(orc, name), err_msg=name)
    s, a, r, s2, d = ring.sample(64)
    assert s.shape == (64, W) and a.shape == (64,) and d.dtype == torch.bool


def test_ring_rejects_bad_input(cuda_device):
    from pulselib_b200.replay import TransitionRing

    ring = TransitionRing(cuda_device, 100, 16)
    z = lambda *s, dt=torch.float32: torch.zeros(*s, dtype=dt, device=cuda_device)
    with pytest.raises(ValueError):
        ring.append(z(200, 16), z(200, dt=torch.int64), z(200), z(200, 16), z(200, dt=torch.bool), z(200, dt=torch.int32))
    with pytest.raises(ValueError):
        ring.append(z(50, 12), z(50, dt=torch.int64), z(50), z(50, 12), z(50, dt=torch.bool), z(50, dt=torch.int32))
