"""GPU suite: the CUDA PokerGPU env replays every env call the reference's OWN unit tests make
(tests/golden/poker_reftests.npz, recorded from /root/reference/tests/poker/test_poker_gpu_*.py while they
passed on the reference; see tests/test_reference_kat.py for the CPU side of the same replay)."""
import numpy as np
import pytest
import torch

from tests import poker_util as pu

pytestmark = pytest.mark.gpu


from tests.kat_dropin import DropinKat


def test_cuda_env_reproduces_reference_unit_tests(cuda_device):
    n = pu.replay_reference_kat(lambda ctor: DropinKat(cuda_device, ctor))
    assert n >= 280
