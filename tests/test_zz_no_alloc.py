"""GPU: pulse_replay_append_static (include/pulse_envs.h) -- the transition hand-off with its scratch in a file-scope
device array of the library, for callers that cannot call cudaMalloc.  It has no counterpart in the reference;
the check is against the existing entry point pulse_replay_append, through
the C ABI, bit for bit (the append moves rows, it computes nothing), on consecutive calls that share the static array.

This file is named so that pytest collects it last: a failure here must not hide
the parity tests."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

RING_FIELDS = ("state", "next", "action", "reward", "done", "total", "appended")


def _ring(dev, cap, W):
    return dict(state=torch.zeros((cap, W), dtype=torch.float32, device=dev),
                next=torch.zeros((cap, W), dtype=torch.float32, device=dev),
                action=torch.zeros(cap, dtype=torch.int64, device=dev),
                reward=torch.zeros(cap, dtype=torch.float32, device=dev),
                done=torch.zeros(cap, dtype=torch.bool, device=dev),
                total=torch.zeros(1, dtype=torch.int64, device=dev),
                appended=torch.zeros(1, dtype=torch.int64, device=dev))


def _step_batch(rng, N, W, dev):
    """one step's tensors, already on the device (the entry points allocate nothing; the caller brings everything)"""
    state = rng.standard_normal((N, W)).astype(np.float32)
    nxt = rng.standard_normal((N, W)).astype(np.float32)
    acts = rng.integers(0, 13, N).astype(np.int64)
    rew = rng.standard_normal(N).astype(np.float32)
    dones = rng.random(N) < 0.15
    seat = rng.integers(0, 3, N).astype(np.int32)
    term = rng.random(N) < 0.3
    return [torch.from_numpy(a).to(dev) for a in (state, nxt, acts, rew, dones, seat, term)]


def _call(lib, static, b, ring, cap, W, q_seat, scratch, dev, use_term=True):
    from pulselib_b200 import _lib
    from pulselib_b200._lib import c_i, ptr

    state, nxt, acts, rew, dones, seat, term = b
    head = (ptr(state), ptr(nxt), c_i(W), ptr(acts), ptr(rew), ptr(dones), ptr(seat), ptr(term if use_term else None),
            c_i(q_seat), ptr(ring["state"]), ptr(ring["next"]), ptr(ring["action"]), ptr(ring["reward"]),
            ptr(ring["done"]), C.c_longlong(cap), ptr(ring["total"]), ptr(ring["appended"]))
    n, stream = C.c_longlong(state.shape[0]), _lib.cuda_stream(dev)
    if static:
        return lib.pulse_replay_append_static(*head, n, stream)
    return lib.pulse_replay_append(*head, ptr(scratch), n, stream)


def _same(a, b, tag):
    for f in RING_FIELDS:
        assert torch.equal(a[f], b[f]), f"{tag}: ring field {f} differs from pulse_replay_append's"


@pytest.mark.parametrize("W", [16, 13])     # vectorised row mover and the 4-byte one
def test_static_scratch_append_matches_append(W, cuda_device):
    from pulselib_b200 import _lib

    lib = _lib.load()
    lib.pulse_replay_scratch_bytes.restype = C.c_size_t
    lib.pulse_replay_static_max_games.restype = C.c_longlong
    assert lib.pulse_replay_static_max_games() == 4194304
    dev = cuda_device
    rng = np.random.default_rng(20261105 + W)
    cap = 140_000
    sizes = [100_003, 257, 70_001, 100_003, 1, 131_072]   # big, small, big again: stale counts beyond nb must not matter
    a, b = _ring(dev, cap, W), _ring(dev, cap, W)
    scratch = torch.empty(int(lib.pulse_replay_scratch_bytes(C.c_longlong(max(sizes)))), dtype=torch.uint8, device=dev)
    batches = [_step_batch(rng, n, W, dev) for n in sizes]
    torch.cuda.synchronize(dev)
    with _lib.device_guard(dev):
        for i, bt in enumerate(batches):
            free_before = torch.cuda.mem_get_info(dev)[0]
            for rep in range(2):                       # two calls in a row on the same inputs, no sync in between
                _lib.check(_call(lib, True, bt, a, cap, W, 1, None, dev, use_term=(i != 3)), "pulse_replay_append_static")
                _lib.check(_call(lib, False, bt, b, cap, W, 1, scratch, dev, use_term=(i != 3)), "pulse_replay_append")
            torch.cuda.synchronize(dev)
            if i > 0:   # the first call loads the kernels (lazy module loading); after that nothing may be allocated
                assert torch.cuda.mem_get_info(dev)[0] >= free_before, "device memory was allocated during the call"
            _same(a, b, f"W={W} batch {i} (n={sizes[i]})")
            assert int(a["total"].item()) > 0
    # the ring has wrapped (capacity 140,000 < rows appended) and both agree on where it stands
    assert int(a["total"].item()) == int(b["total"].item()) > cap


def test_static_scratch_append_alone_after_other_work(cuda_device):
    """Back-to-back calls with different selections, each checked: the second call must not see the first one's counts."""
    from pulselib_b200 import _lib

    lib = _lib.load()
    lib.pulse_replay_scratch_bytes.restype = C.c_size_t
    dev, W, cap, N = cuda_device, 16, 50_000, 40_000
    rng = np.random.default_rng(7)
    a, b = _ring(dev, cap, W), _ring(dev, cap, W)
    scratch = torch.empty(int(lib.pulse_replay_scratch_bytes(C.c_longlong(N))), dtype=torch.uint8, device=dev)
    with _lib.device_guard(dev):
        for q in (0, 1, 2, 0):
            bt = _step_batch(rng, N, W, dev)
            _lib.check(_call(lib, True, bt, a, cap, W, q, None, dev), "pulse_replay_append_static")
            _lib.check(_call(lib, False, bt, b, cap, W, q, scratch, dev), "pulse_replay_append")
            seat, term = bt[5], bt[6]
            want = int(((seat == q) & ~term).sum().item())
            assert int(a["appended"].item()) == want == int(b["appended"].item())
            _same(a, b, f"q_seat={q}")


def test_static_scratch_append_rejects_oversize(cuda_device):
    from pulselib_b200 import _lib
    from pulselib_b200._lib import c_i, ptr

    lib = _lib.load()
    dev = cuda_device
    t = torch.zeros(16, dtype=torch.float32, device=dev)   # never touched: the size check comes first
    args = (ptr(t), ptr(t), c_i(16), ptr(t), ptr(t), ptr(t), ptr(t), ptr(None), c_i(0), ptr(t), ptr(t), ptr(t), ptr(t),
            ptr(t), C.c_longlong(1 << 40), ptr(t), ptr(t))
    with _lib.device_guard(dev):
        rc = lib.pulse_replay_append_static(*args, C.c_longlong(4194304 + 1), _lib.cuda_stream(dev))
    assert rc != 0
    msg = lib.pulse_last_error().decode()
    assert "4194304" in msg and "exceeds" in msg, msg
    with pytest.raises(_lib.PulseError):
        _lib.check(rc, "pulse_replay_append_static")
    torch.cuda.synchronize(dev)


def test_static_scratch_append_at_the_limit(cuda_device):
    """n_games = the stated maximum: 16,384 block counts plus the base word fill the static array exactly as laid out."""
    from pulselib_b200 import _lib

    lib = _lib.load()
    lib.pulse_replay_scratch_bytes.restype = C.c_size_t
    dev, W, N = cuda_device, 16, 4194304
    cap = N
    g = torch.Generator(device=dev).manual_seed(11)
    bt = [torch.randn((N, W), generator=g, device=dev), torch.randn((N, W), generator=g, device=dev),
          torch.randint(0, 13, (N,), generator=g, device=dev), torch.randn(N, generator=g, device=dev),
          torch.rand(N, generator=g, device=dev) < 0.15, torch.randint(0, 2, (N,), generator=g, device=dev).to(torch.int32),
          torch.rand(N, generator=g, device=dev) < 0.3]
    a, b = _ring(dev, cap, W), _ring(dev, cap, W)
    scratch = torch.empty(int(lib.pulse_replay_scratch_bytes(C.c_longlong(N))), dtype=torch.uint8, device=dev)
    with _lib.device_guard(dev):
        for rep in range(2):
            _lib.check(_call(lib, True, bt, a, cap, W, 1, None, dev), "pulse_replay_append_static")
            _lib.check(_call(lib, False, bt, b, cap, W, 1, scratch, dev), "pulse_replay_append")
            torch.cuda.synchronize(dev)
            _same(a, b, f"limit, call {rep}")
    assert int(a["appended"].item()) == int(((bt[5] == 1) & ~bt[6]).sum().item())
