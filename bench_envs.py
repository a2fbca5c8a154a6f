"""Throughput of the non-poker env steps (BASELINE.json configs: 2048 1M boards, Tetris 2M envs, Particle2D
16M envs, Blackjack 4096 envs), used by bench.py for `--workload <name>` and for the compact `workloads`
block of the default line.  Same rules as bench.py: CUDA events on the launching stream, >= 3 warm-up
steps, inputs larger than L2 between touches (small configs rotate over several independent env batches),
one launch per env step, K steps captured per CUDA graph where a step is shorter than the host's
per-call overhead.
"""
from __future__ import annotations

import os
import time

SEED = 20260401
L2_BYTES = 126e6

SPECS = {
    # algo_bytes: SURVEY.md 8(d) "algorithmic bytes per env-step" (fixed there; DESIGN.md 4.x derives them)
    "tfe_4x4": dict(n=1 << 20, algo_bytes=106, replicas=4, kernel="tfe_step_kernel<4, uint8_t>", graph=32, actions=16,
                    desc="2048 4x4, {n} boards/GPU, uniform random moves (uint8), spawns from the chance tape, "
                         "auto-reset on game over, int32 4x4 observation written every step"),
    # open-loop action tapes, 16 steps per launch (pulse_tfe_rollout): the row table is staged once per 16 steps and a
    # board stays in registers between its steps -- action 1 + obs 64 + reward 4 + done 1 = 70 B per board-step plus
    # 32 B of state per 16 steps; reported against the same fixed 106 B
    "tfe_4x4_rollout": dict(n=1 << 20, algo_bytes=106, moved_bytes=72, replicas=4, kernel="tfe_rollout_kernel<4, uint8_t>",
                            graph=0, actions=16, steps_per_call=16,
                            desc="2048 4x4, {n} boards/GPU, uniform random moves (uint8) supplied as a [16, N] tape, 16 steps per "
                                 "launch, spawns from the chance tape, auto-reset on game over, int32 4x4 observations of "
                                 "every step written ([16, N, 4, 4])"),
    # moved_bytes: what the kernel can actually move per env-step -- the rows tensor IS the observation (no 40 B copy):
    # rows 40 in + 40 out, piece 1 + 1, lines 4 + 4, done 1 + 1, action 1, reward 4 = 97 B; reported next to the fixed 134 B
    "tetris": dict(n=1 << 21, algo_bytes=134, moved_bytes=97, replicas=4, kernel="tetris_step_kernel<uint8_t>", graph=32, actions=16,
                   desc="Tetris 10x20, {n} envs/GPU, uniform random placements over 40 actions (uint8), pieces from "
                        "the chance tape, auto-reset on lock-out, rows tensor is the observation"),
    "particle2d": dict(n=1 << 24, algo_bytes=70, kernel="particle2d_step_kernel", graph=0, actions=4,
                       desc="Particle2D fp32, {n} envs/GPU, dt=0.1, max_steps=200, uniform actions in [-1,1]^2"),
    # build_actions + step as ONE launch (scripted seats evaluated inside the step kernel),
    # per-table auto-reset.  Bytes: the step's 201 B with the 8 B action read replaced by the 8 B write of
    # the action that was played.
    "poker_hu_scripted": dict(n=1 << 22, algo_bytes=201, kernel="poker_packed_step_kernel<2, 2, 8, true, 2>", graph=0,
                              actions=1,
                              desc="PokerGPU heads-up, {n} tables/GPU, seats played by the scripted TightAggressive and "
                                   "SmallBall rules inside the step kernel (no action tensor), draws from the chance "
                                   "tape, per-table auto-reset"),
    # SURVEY 8(b) pulse_poker_rollout, "on-device random policy, K steps per launch": one call deals the whole batch and
    # plays a 16-step hand with every seat's action drawn on the device (PULSE_AGENT_RANDOM, chance tape) -- the
    # headline's workload without an action tensor and with one launch per hand instead of sixteen.  Written per step:
    # reward + done of every table ([16, N]); the observation only after the last step.
    "poker_hu_rollout": dict(n=1 << 22, algo_bytes=201, kernel="poker_packed_rollout_kernel<2> (+ reset)", graph=0, actions=16,
                             steps_per_call=16,
                             desc="PokerGPU heads-up, {n} tables/GPU, uniform random actions 0..12 drawn ON THE DEVICE from the "
                                  "chance tape, 16 steps per launch, whole-batch reset before every launch (inside the timed "
                                  "region), rewards / dones of every step and the final observation written"),
    # the headline's own trajectories (supplied uniform random actions), 16 steps per launch: pulse_poker_packed_rollout_actions
    "poker_hu_rollout_tape": dict(n=1 << 22, algo_bytes=201, kernel="poker_packed_rollout_kernel<2> (+ reset)", graph=0, actions=16,
                                  steps_per_call=16,
                                  desc="PokerGPU heads-up, {n} tables/GPU, uniform random actions 0..12 supplied as a uint8 [16, N] "
                                       "tape, 16 steps per launch, whole-batch reset before every launch (inside the timed region), "
                                       "rewards / dones of every step and the final observation written"),
    # SURVEY 8(d) config 3, convention (ii): supplied uniform random actions, per-table auto-reset -- no table ever
    # idles; a table that finished is dealt its next hand by the step that follows (that step plays no action for it).
    "poker_hu_autoreset": dict(n=1 << 22, algo_bytes=201, kernel="poker_packed_step_kernel<2, 2, 10, false, 2> + redeal pass",
                               graph=0, actions=16,
                               desc="PokerGPU heads-up, {n} tables/GPU, uniform random actions 0..12 supplied (int64), "
                                    "per-table auto-reset: every table-step is live or deals the table's next hand"),
    # SURVEY 8(f) N1+N2+N3 together: what one iteration of the reference's training loop does around the step
    # (scripts/Poker/trainGPU_performance.py:176-206) with everything on the device: scripted opponent inside the
    # step, learning seat's actions supplied, hand statistics folded, learning-seat transitions appended to the
    # ring; state/next_state ping-pong instead of state.clone().  Bytes: step 201 + stats 2 + ring select 5 +
    # ~half the tables' transitions (2 rows read + written, 141 B each way).
    "poker_hu_train_loop": dict(n=1 << 22, algo_bytes=201 + 2 + 5 + 141, kernel="step<2,2,8,true,2> + redeal + hand_stats + replay x3",
                                graph=0, actions=16,
                                desc="PokerGPU heads-up, {n} tables/GPU, seat 0 = learner (uniform random actions supplied), "
                                     "seat 1 = TightAggressive rule inside the step kernel, per-table auto-reset, hand "
                                     "statistics and the learner's transitions (ring of 8M) gathered on the device every step"),
    "blackjack": dict(n=4096, algo_bytes=50, kernel="blackjack_rollout_kernel<unsigned char>", graph=0, actions=128, steps_per_call=128,
                      desc="Blackjack, {n} envs/GPU (the reference's CPU-sized config), uniform hit/stand, hands of 8 steps: "
                           "ONE launch deals and plays 16 hands (pulse_blackjack_rollout, 128 steps, outputs [128, N, ...]); "
                           "working set is L2-resident by construction"),
    # the same config one launch per step (the reference's call pattern, blackjack.py:179-186)
    "blackjack_per_step": dict(n=4096, algo_bytes=50, kernel="blackjack_step_kernel", graph=0, actions=8,
                               desc="Blackjack, {n} envs/GPU, uniform hit/stand, one launch per step, whole-batch reset every "
                                    "8 steps inside the timed region"),
}


def _make(name, dev, n, rank, e2e=False):
    """-> (env, actions [L, n, ...] on the device, step(env, a) -> tuple of output tensors)"""
    import torch

    spec = SPECS[name]
    gen = torch.Generator(device=dev)
    gen.manual_seed(SEED + 17 * rank + (1 if e2e else 0))
    L = spec["actions"]
    if name == "tfe_4x4":
        from pulselib_b200.envs.tfe import TFEBatched

        env = TFEBatched(dev, n, side=4, seed=SEED, env_base=rank * n, auto_reset=True, device_counter=True)
        acts = torch.randint(0, 4, (L, n), dtype=torch.uint8, device=dev, generator=gen)

        def step(e, a):
            o, r, d, _, _ = e.step(a)
            return o, r, d
    elif name == "tfe_4x4_rollout":
        from pulselib_b200.envs.tfe import TFEBatched

        env = TFEBatched(dev, n, side=4, seed=SEED, env_base=rank * n, auto_reset=True)
        acts = torch.randint(0, 4, (L, n), dtype=torch.uint8, device=dev, generator=gen)
        env._bench_k, env._bench_out = 0, None

        def step(e, a):                                  # one launch plays the next L steps of every board
            o, r, d = e.rollout(acts)
            return o[-1], r[-1], d[-1]
    elif name == "tetris":
        from pulselib_b200.envs.tetris import TetrisBatched

        env = TetrisBatched(dev, n, seed=SEED, env_base=rank * n, auto_reset=True, device_counter=True)
        acts = torch.randint(0, 40, (L, n), dtype=torch.uint8, device=dev, generator=gen)

        def step(e, a):
            o, r, d, _, _ = e.step(a)
            return o, r, d
    elif name == "particle2d":
        from pulselib_b200.envs.particle2d import Particle2D

        env = Particle2D(dev, n, dt=0.1, max_steps=200, seed=SEED, env_base=rank * n)
        acts = torch.rand((L, n, 2), dtype=torch.float32, device=dev, generator=gen) * 2 - 1

        def step(e, a):
            o, r, t, _, _ = e.step(a)
            return o, r, t
    elif name == "poker_hu_scripted":
        from pulselib_b200 import policy
        from pulselib_b200.envs.poker_packed import PokerGPUPacked

        env = PokerGPUPacked(dev, n_players=2, max_players=2, n_games=n, starting_bbs=100, max_bbs=1000, w1=.5, w2=.3,
                             K=100, alpha=50, seed=SEED, env_base=rank * n, auto_reset=True)
        acts = torch.zeros((L, 1), dtype=torch.int64, device=dev)
        kinds = bytes([policy.TIGHT_AGGRESSIVE, policy.SMALL_BALL])

        def step(e, a):
            o, r, d, _, info = e.step_policy(kinds)
            return o, r, d, info["actions"]
    elif name == "poker_hu_rollout":
        from pulselib_b200 import policy
        from pulselib_b200.envs.poker_packed import PokerGPUPacked

        env = PokerGPUPacked(dev, n_players=2, max_players=2, n_games=n, starting_bbs=100, max_bbs=1000, w1=.5, w2=.3,
                             K=100, alpha=50, seed=SEED, env_base=rank * n, auto_reset=False, live_info=False)
        acts = torch.zeros((L, 1), dtype=torch.int64, device=dev)
        kinds = bytes([policy.RANDOM, policy.RANDOM])

        def step(e, a):                                  # one hand: deal + one launch of 16 steps
            e.reset()
            o, r, d, _ = e.rollout_policy(kinds, L, return_actions=False)
            return o, r[-1], d[-1]
    elif name == "poker_hu_rollout_tape":
        from pulselib_b200.envs.poker_packed import PokerGPUPacked

        env = PokerGPUPacked(dev, n_players=2, max_players=2, n_games=n, starting_bbs=100, max_bbs=1000, w1=.5, w2=.3,
                             K=100, alpha=50, seed=SEED, env_base=rank * n, auto_reset=False, live_info=False)
        acts = torch.randint(0, 13, (L, n), dtype=torch.uint8, device=dev, generator=gen)

        def step(e, a):                                  # one hand: deal + one launch of 16 steps on the tape
            e.reset()
            o, r, d = e.rollout_actions(acts)
            return o, r[-1], d[-1]
    elif name == "poker_hu_autoreset":
        from pulselib_b200.envs.poker_packed import PokerGPUPacked

        env = PokerGPUPacked(dev, n_players=2, max_players=2, n_games=n, starting_bbs=100, max_bbs=1000, w1=.5, w2=.3,
                             K=100, alpha=50, seed=SEED, env_base=rank * n, auto_reset=True, live_info=False)
        acts = torch.randint(0, 13, (L, n), dtype=torch.int64, device=dev, generator=gen)
        env._rew = torch.empty(n, dtype=torch.float32, device=dev)
        env._done = torch.empty(n, dtype=torch.bool, device=dev)

        def step(e, a):
            e.step_into(a, e.obs, e._rew, e._done)
            return e.obs, e._rew, e._done
    elif name == "poker_hu_train_loop":
        from pulselib_b200 import policy
        from pulselib_b200.envs.poker_packed import PokerGPUPacked
        from pulselib_b200.hand_stats import HandStats
        from pulselib_b200.replay import TransitionRing

        env = PokerGPUPacked(dev, n_players=2, max_players=2, n_games=n, starting_bbs=100, max_bbs=1000, w1=.5, w2=.3,
                             K=100, alpha=50, seed=SEED, env_base=rank * n, auto_reset=True)
        acts = torch.randint(0, 13, (L, n), dtype=torch.int64, device=dev, generator=gen)
        kinds = bytes([policy.EXTERNAL, policy.TIGHT_AGGRESSIVE])
        env._hs = HandStats(env, q_seat=0)
        env._ring = TransitionRing(dev, 2 * n, env.obs_size)
        env._pp = [torch.zeros_like(env.obs), torch.zeros_like(env.obs)]
        env._k = 0

        def step(e, a):
            cur, nxt = e._pp[e._k & 1], e._pp[(e._k + 1) & 1]
            o, r, d, _, info = e.step_policy(kinds, actions=a, obs_out=nxt, return_seat=True)
            e._hs.update(d)                      # auto-reset: every done flag is one finished hand (rearm mode)
            e._ring.append(cur, info["actions"], r, nxt, d, info["seat_idx_before"], None, 0)
            e._k += 1
            return o, r, d
    elif name == "blackjack":
        from pulselib_b200.envs.blackjack import BlackJack

        env = BlackJack(dev, n, seed=SEED, env_base=rank * n)
        acts = torch.randint(0, 2, (L, n), dtype=torch.uint8, device=dev, generator=gen)
        env._bench_k, env._bench_out = 0, None

        def step(e, a):                                  # one launch: 16 x (deal + the hand's 8 steps)
            o, r, t = e.rollout(acts, reset_first=True, n_hands=L // 8)[:3]
            return o[-1], r[-1], t[-1]
    elif name == "blackjack_per_step":
        from pulselib_b200.envs.blackjack import BlackJack

        env = BlackJack(dev, n, seed=SEED, env_base=rank * n)
        acts = torch.randint(0, 2, (L, n), dtype=torch.int64, device=dev, generator=gen)
        env._bench_k = 0

        def step(e, a):
            if e._bench_k % 8 == 0:
                e.reset()
            e._bench_k += 1
            o, r, t, _, _ = e.step(a)
            return o, r, t
    else:
        raise KeyError(name)
    first_obs = env.reset()
    if name == "poker_hu_train_loop":
        env._pp[0].copy_(first_obs[0])
        env._hs.begin_hand()
    return env, acts, step


def measure(name, dev, rank, world, steps, warmup, n=None, dist=None):
    """Device-timed throughput of `steps` env steps; returns a dict (value is per GPU)."""
    import torch

    spec = SPECS[name]
    n = int(n or spec["n"])
    per_step_bytes = spec["algo_bytes"] * n
    # independent env batches stepped in rotation, so a batch's state was evicted from the 126 MB L2 before it is
    # touched again (explicit where the state is much smaller than the algorithmic bytes: Tetris rows are 84 MB)
    replicas = 1 if name.startswith("blackjack") else max(spec.get("replicas", 1), int(-(-2.5 * L2_BYTES // per_step_bytes)))
    made = [_make(name, dev, n, rank * 64 + r) for r in range(replicas)]
    envs, acts, step = [m[0] for m in made], made[0][1], made[0][2]
    L = acts.shape[0]
    stream = torch.cuda.current_stream(dev)
    launches = 0
    chunk = spec["graph"]
    if chunk:
        chunk = min(chunk, steps)
        chunk -= chunk % replicas or 0
        chunk = max(chunk, replicas)

    spc = int(spec.get("steps_per_call", 1))   # rollout workloads: one call (= one launch) plays this many steps
    EVENT_EVERY = max(1, int(os.environ.get("PULSE_BENCH_EVENT_EVERY", "5"))) if spc == 1 else 1

    def eager(k, start, pairs=None):
        for n_call, i in enumerate(range(0, k, spc)):
            j = start + i
            timed = pairs is not None and n_call % EVENT_EVERY == 0   # a sample of the launches (see bench.py)
            if timed:
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
            step(envs[j % replicas], acts[j % L])
            if timed:
                e1.record(stream)
                pairs.append((e0, e1))

    graph = None
    if chunk:
        side = torch.cuda.Stream(dev)
        side.wait_stream(stream)
        with torch.cuda.stream(side):
            eager(max(warmup, replicas), 0)
        stream.wait_stream(side)
        torch.cuda.synchronize(dev)
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            eager(chunk, 0)
        stream = torch.cuda.current_stream(dev)
        graph.replay()                                   # warm replay
    else:
        eager(max(warmup, 3), 0)

    def barrier():
        torch.cuda.synchronize(dev)
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize(dev)

    live_fraction = None
    if name in ("poker_hu_autoreset", "poker_hu_scripted"):
        # tables that finish on a step are dealt by the next one instead of playing: live share of the table-steps
        fin = 0.0
        for j in range(32):
            out = step(envs[0], acts[j % L])
            fin += float(out[2].float().mean())
        live_fraction = 1.0 - fin / 32

    barrier()
    pairs = []
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record(stream)
    if graph is not None:
        reps, rem = divmod(steps, chunk)
        for _ in range(reps):
            graph.replay()
        eager(rem, 0)
    else:
        eager(steps, 0, pairs)
    t1.record(stream)
    barrier()
    total_ms = t0.elapsed_time(t1)
    kernel_ms = (sum(a.elapsed_time(b) for a, b in pairs) / len(pairs) / spc) if pairs else total_ms / steps
    launches = steps // 16 + 1 if name == "tfe_4x4_rollout" else 2 * (steps // 16 + 1) if name in ("poker_hu_rollout", "poker_hu_rollout_tape") else steps // 128 + 1 if name == "blackjack" else steps + (steps // 8 + 1 if name == "blackjack_per_step" else 0)
    return dict(n=n, total_ms=total_ms, kernel_ms=kernel_ms, launches=launches, replicas=replicas, live_fraction=live_fraction,
                mode=f"cuda graph of {chunk} steps" if graph is not None else "eager launches",
                kernel_time="graph time / steps (includes inter-kernel gaps)" if graph is not None
                else ("CUDA events around every %dth step launch of the timed region" % EVENT_EVERY if EVENT_EVERY > 1 else "CUDA events around each step launch"),
                value=n * steps / (total_ms * 1e-3))


def e2e(name, dev, rank, steps, n=None, chunks=4):
    """Same metric through the public env API with HOST buffers: pinned actions -> device, env.step, outputs
    -> pinned host, every step; `chunks` independent env batches on their own streams so copies overlap."""
    import torch

    spec = SPECS[name]
    n = int(n or spec["n"])
    if name == "blackjack":
        return _e2e_blackjack_rollout(dev, rank, steps, n)
    if name in ("tfe_4x4", "tfe_4x4_rollout"):
        return _e2e_tfe_host(dev, rank, steps, n, min(chunks, 4))   # 1M boards: three to six chunks measure the same, eight lose 10%
    chunks = chunks if n >= 1 << 16 else 1
    nc = n // chunks
    made = [_make(name, dev, nc, rank * 64 + 32 + c, e2e=True) for c in range(chunks)]
    step = made[0][2]
    streams = [torch.cuda.Stream(dev) for _ in range(chunks)]
    h_act = [m[1].cpu().pin_memory() for m in made]
    d_act = [torch.empty_like(m[1][0]) for m in made]
    outs = [step(m[0], m[1][0]) for m in made]
    h_out = [[torch.empty(o.shape, dtype=o.dtype).pin_memory() for o in out] for out in outs]
    L = h_act[0].shape[0]
    torch.cuda.synchronize(dev)

    def run(k):
        for i in range(k):
            for c in range(chunks):
                with torch.cuda.stream(streams[c]):
                    d_act[c].copy_(h_act[c][i % L], non_blocking=True)
                    out = step(made[c][0], d_act[c])
                    for h, o in zip(h_out[c], out):
                        h.copy_(o, non_blocking=True)

    run(3)
    torch.cuda.synchronize(dev)
    w0 = time.perf_counter()
    run(steps)
    for s in streams:
        s.synchronize()
    ms = (time.perf_counter() - w0) * 1e3
    h2d = sum(a[0].numel() * a.element_size() for a in h_act)
    d2h = sum(h.numel() * h.element_size() for hs in h_out for h in hs)
    return dict(value=nc * chunks * steps / (ms * 1e-3), ms=ms, steps=steps, chunks=chunks, h2d_bytes_per_step=h2d,
                d2h_bytes_per_step=d2h)


def _e2e_tfe_host(dev, rank, steps, n, chunks):
    """2048 end to end through the host-buffer env (pulselib_b200.host_env.TFEHostEnv -> pulse_tfe_host_step): numpy uint8
    moves in, int32 tile values [N, 4, 4] + int32 rewards + bool dones out in host arrays every step.  1 B per board goes
    to the GPU, the packed 8 B board + 5 B come back, host threads rebuild the tile values.  Host-synchronous API: wall
    clock around the loop."""
    import numpy as np
    import torch

    from pulselib_b200.host_env import TFEHostEnv

    avail = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    threads = max(1, avail // max(1, int(os.environ.get("LOCAL_WORLD_SIZE", "1"))))
    env = TFEHostEnv(dev, n_envs=n, side=4, seed=SEED, env_base=rank * n, auto_reset=True, chunks=chunks, threads=threads)
    acts = np.random.default_rng(SEED + rank).integers(0, 4, (16, n)).astype(np.uint8)
    env.reset()
    for i in range(3):
        env.step(acts[i])
    torch.cuda.synchronize(dev)
    w0 = time.perf_counter()
    for i in range(steps):
        env.step(acts[i % 16])
    torch.cuda.synchronize(dev)
    ms = (time.perf_counter() - w0) * 1e3
    inf = env.info()
    env.close()
    return dict(value=n * steps / (ms * 1e-3), ms=ms, steps=steps, chunks=inf["chunks"], host_threads=inf["threads"],
                h2d_bytes_per_step=inf["last_h2d_bytes"], d2h_bytes_per_step=inf["last_d2h_bytes"],
                api="pulselib_b200.host_env.TFEHostEnv.step (pulse_tfe_host_step): host arrays in and out, packed boards over PCIe")


def _e2e_blackjack_rollout(dev, rank, steps, n):
    """Hands of 8 steps with HOST buffers: the hand's [8, N] action tape goes to the device, one launch deals and plays
    the hand, the [8, N, ...] outputs come back -- per hand, inside the timed region."""
    import torch
    from pulselib_b200.envs.blackjack import BlackJack

    env = BlackJack(dev, n, seed=SEED, env_base=rank * n)
    KT = 128                                         # 16 hands of 8 steps per launch
    hands = max(16, steps // KT)                     # launches timed
    h_act = torch.randint(0, 2, (KT, n), dtype=torch.uint8).pin_memory()
    d_act = torch.empty((KT, n), dtype=torch.uint8, device=dev)
    h_obs = torch.empty((KT, n, 3), dtype=torch.int32).pin_memory()
    h_rew = torch.empty((KT, n), dtype=torch.int32).pin_memory()
    h_term = torch.empty((KT, n), dtype=torch.bool).pin_memory()

    def run(k):
        for _ in range(k):
            d_act.copy_(h_act, non_blocking=True)
            o, r, t, _ = env.rollout(d_act, reset_first=True, n_hands=KT // 8)
            h_obs.copy_(o, non_blocking=True); h_rew.copy_(r, non_blocking=True); h_term.copy_(t, non_blocking=True)
        torch.cuda.synchronize(dev)

    run(3)
    w0 = time.perf_counter()
    run(hands)
    ms = (time.perf_counter() - w0) * 1e3
    return dict(value=n * KT * hands / (ms * 1e-3), ms=ms, steps=KT * hands, chunks=1, h2d_bytes_per_step=n,
                d2h_bytes_per_step=n * (12 + 4 + 1))


def cpu_baseline(name, min_seconds=10.0):
    """The oracle port of the reference step (oracle/*.c, OpenMP over envs) on this box's host cores, on a
    bounded sample of the same workload.  Chance words / pieces are pre-generated outside the timed region."""
    import ctypes as C

    import numpy as np
    from oracle import pyoracle

    avail = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    cores = int(pyoracle.lib().oracle_set_threads(avail))   # torchrun exports OMP_NUM_THREADS=1: set it explicitly
    rng = np.random.default_rng(SEED)
    spent, steps = 0.0, 0
    if name in ("tfe_4x4", "tfe_4x4_rollout"):
        n = 1 << 20
        orc = pyoracle.OracleTFE(n, 4)
        orc.reset(rng.integers(0, 2**32, (n, 4), dtype=np.uint64).astype(np.uint32))
        ch = [rng.integers(0, 2**32, (n, 2), dtype=np.uint64).astype(np.uint32) for _ in range(4)]
        rc = rng.integers(0, 2**32, (n, 4), dtype=np.uint64).astype(np.uint32)
        acts = rng.integers(0, 4, (8, n)).astype(np.int64)
        k = 0
        while spent < min_seconds:
            t0 = time.perf_counter()
            orc.step(acts[k % 8], ch[k % 4], auto_reset=True, reset_chance4=rc)
            spent += time.perf_counter() - t0
            steps += n; k += 1
        sample = f"{n} boards x {k} steps, oracle/oracle_tfe.c (int32 4x4 boards, the reference's representation)"
    elif name == "tetris":
        n = 1 << 20
        orc = pyoracle.OracleTetris(n)
        orc.reset(rng.integers(0, 7, n).astype(np.uint8))
        pcs = [rng.integers(0, 7, n).astype(np.uint8) for _ in range(4)]
        acts = rng.integers(0, 40, (8, n)).astype(np.int64)
        k = 0
        while spent < min_seconds:
            t0 = time.perf_counter()
            orc.step(acts[k % 8], pcs[k % 4], auto_reset=True, reset_piece=pcs[(k + 1) % 4])
            spent += time.perf_counter() - t0
            steps += n; k += 1
        sample = f"{n} envs x {k} steps, oracle/oracle_tetris.c (this repo's spec; the reference has no Tetris logic)"
    elif name == "particle2d":
        n = 1 << 22
        L = pyoracle.lib()
        state = np.zeros((n, 4), np.float32)
        state[:, :2] = rng.standard_normal((n, 2)).astype(np.float32) * 5
        cnt, rew, term = np.zeros(n, np.int32), np.zeros(n, np.float32), np.zeros(n, np.uint8)
        acts = rng.uniform(-1, 1, (4, n, 2)).astype(np.float32)
        k = 0
        while spent < min_seconds:
            t0 = time.perf_counter()
            L.oracle_particle2d_step(pyoracle.p(state), pyoracle.p(acts[k % 4]), pyoracle.p(cnt), pyoracle.p(rew),
                                     pyoracle.p(term), C.c_float(0.1), C.c_int(200), C.c_int(n))
            spent += time.perf_counter() - t0
            steps += n; k += 1
        sample = f"{n} envs x {k} steps, oracle/oracle_core.c::oracle_particle2d_step"
    elif name in ("poker_hu_scripted", "poker_hu_autoreset"):
        from pulselib_b200 import handranks

        n = 1 << 20
        kw = dict(n_players=2, max_players=2, n_games=n, starting_bbs=100, max_bbs=1000, w1=.5, w2=.3, K=100, alpha=50)
        orc = pyoracle.OraclePoker(handranks.load_table(), **kw)
        kinds = np.array([3, 5], np.uint8)
        ch = [rng.integers(0, 2**32, (n, 2), dtype=np.uint64).astype(np.uint32) for _ in range(4)]
        acts = np.zeros(n, np.int64)
        k = hands = 0
        while spent < min_seconds:
            orc.tape_decks(SEED, 0, hands)
            t0 = time.perf_counter()
            orc.reset(None)
            for i in range(16):   # whole-batch hands: the oracle has no per-table auto-reset
                pyoracle.poker_policy(orc.obs, orc.idx, kinds, ch[i % 4], acts)
                orc.step(acts)
            spent += time.perf_counter() - t0
            steps += 16 * n; k += 16; hands += 1
        sample = f"{n} tables x {hands} hands x 16 steps (oracle policy + reset + step, obs built), oracle/oracle_poker.c"
    elif name == "poker_hu_train_loop":
        from pulselib_b200 import handranks

        n = 1 << 20
        kw = dict(n_players=2, max_players=2, n_games=n, starting_bbs=100, max_bbs=1000, w1=.5, w2=.3, K=100, alpha=50)
        orc = pyoracle.OraclePoker(handranks.load_table(), **kw)
        ring = pyoracle.OracleRing(2 * n, orc.obs_size)
        kinds = np.array([0, 3], np.uint8)
        ch = [rng.integers(0, 2**32, (n, 2), dtype=np.uint64).astype(np.uint32) for _ in range(4)]
        ext = rng.integers(0, 13, (4, n)).astype(np.int64)
        vec = np.zeros(67)
        k = hands = 0
        while spent < min_seconds:
            orc.tape_decks(SEED, 0, hands)
            t0 = time.perf_counter()
            state = orc.reset(None).copy()
            initial, term = orc.stacks[:, 0].copy(), np.zeros(n, bool)
            for i in range(16):
                acts = pyoracle.poker_policy(state, orc.idx, kinds, ch[i % 4], ext[i % 4].copy())
                seat = orc.idx.copy()
                nxt, rew, done = orc.step(acts)
                pyoracle.hand_stats_update(vec, orc.stacks[:, 0], initial, orc.stages, orc.button, done, term.copy(), 0, 2)
                ring.append(state, acts, rew, nxt, done, seat, term, 0)
                term |= done
                state = nxt.copy()
            spent += time.perf_counter() - t0
            steps += 16 * n; k += 16; hands += 1
        sample = (f"{n} tables x {hands} hands x 16 steps: oracle policy + step (OpenMP) + numpy hand statistics and ring "
                  f"append (the reference does these with torch masks), whole-batch hands")
    elif name in ("blackjack", "blackjack_per_step"):
        n = 4096
        orc = pyoracle.OracleBlackjack(n)
        acts = rng.integers(0, 2, (8, n)).astype(np.int64)
        decks = np.argsort(rng.random((n, 52)), axis=1).astype(np.int32)
        k = 0
        while spent < min(min_seconds, 5.0):
            t0 = time.perf_counter()
            if k % 8 == 0:
                orc.reset(decks)
            orc.step(acts[k % 8])
            spent += time.perf_counter() - t0
            steps += n; k += 1
        sample = f"{n} envs x {k} steps (reset every 8), oracle/oracle_blackjack.c, below the OpenMP threshold: 1 thread"
        cores = 1
    else:
        raise KeyError(name)
    return {"value": steps / spent, "unit": "env-steps/s", "cores": cores, "kind": "port",
            "sample": sample + f", OpenMP on {cores} thread(s); the reference itself is Python (torch eager / numba) "
                               f"and cannot travel to this box"}, spent, steps
