"""CPU suite for the transition hand-off row: the numpy restatement of the reference's masked
gathers + ReplayBufferTorch.add against a trace recorded from the reference's own buffer class
(tests/golden/poker_replay.npz, made by tests/golden/make_replay_golden.py)."""
import numpy as np
import pytest

from oracle import pyoracle
from tests import poker_util as pu

Z = np.load(pu.GOLDEN / "poker_replay.npz")


def replay_trace(tag, ring_append):
    N, W, P, q_seat, steps, cap = (int(v) for v in Z[f"{tag}/meta"])
    for t in range(steps):
        ring_append(*(Z[f"{tag}/{k}"][t] for k in ("state", "actions", "rewards", "next_state", "dones", "seat", "term_before")),
                    q_seat)
    return cap, W


def check_ring(tag, state_buf, next_buf, action_buf, reward_buf, done_buf, position, size):
    pos, sz = (int(v) for v in Z[f"{tag}/position_size"])
    assert (position, size) == (pos, sz)
    # the reference stores int32 (truncation toward zero) -- compare after the same cast
    np.testing.assert_array_equal(np.trunc(state_buf).astype(np.int32), Z[f"{tag}/state_buffer"])
    np.testing.assert_array_equal(np.trunc(next_buf).astype(np.int32), Z[f"{tag}/next_state_buffer"])
    np.testing.assert_array_equal(action_buf.astype(np.int32), Z[f"{tag}/action_buffer"])
    np.testing.assert_array_equal(np.trunc(reward_buf).astype(np.int32), Z[f"{tag}/reward_buffer"])
    np.testing.assert_array_equal(done_buf.astype(np.int32), Z[f"{tag}/terminated_buffer"])


@pytest.mark.parametrize("tag", ["wrap", "six"])
def test_oracle_ring_matches_reference_buffer(tag):
    N, W, P, q_seat, steps, cap = (int(v) for v in Z[f"{tag}/meta"])
    ring = pyoracle.OracleRing(cap, W)
    replay_trace(tag, ring.append)
    check_ring(tag, ring.state_buffer, ring.next_state_buffer, ring.action_buffer, ring.reward_buffer,
               ring.terminated_buffer, ring.position, ring.size)
