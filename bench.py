#!/usr/bin/env python
"""Headline benchmark: env-steps/s of the PokerGPU heads-up NLHE table step (BASELINE.json configs[2]:
4M parallel tables, uniform random actions over 0..12, full showdown evaluation), weak-scaled over N
GPUs with 4M tables per GPU and no data-path collective (tables are independent).

A "step" is one `env.step(actions)` over the whole batch.  Hands follow the reference's convention
(scripts/Poker/trainGPU.py:79-108): whole-batch reset, then HAND_LEN steps, every table counted at
every step; the reset launches sit inside the timed region.  One JSON line is printed by rank 0.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload poker_hu|poker_6max|poker_10max|...]
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

HAND_LEN = 16          # steps per whole-batch hand (max observed heads-up hand is 14)
SEED = 20260401        # scripts/Poker/trainGPU_performance.py:52
CFG = dict(starting_bbs=100, max_bbs=1000, w1=.5, w2=.3, K=100, alpha=50)  # config/pokerGPU.yaml:26-33
WORKLOADS = {
    "poker_hu": dict(n_players=2, max_players=2, tables=4_194_304, algo_bytes=201, hand_len=16),
    "poker_6max": dict(n_players=6, max_players=6, tables=4_194_304, algo_bytes=401, hand_len=24),
    # the reference's own training default (PokerGPU.py:20, scripts/Poker/trainGPU.py:177-188)
    "poker_10max": dict(n_players=10, max_players=10, tables=4_194_304, algo_bytes=601, hand_len=32),
    # the same heads-up workload through the registered drop-in: gym.make("Pulse-Poker-GPU-v1") and the env-facing
    # part of the reference's loop body (scripts/Poker/trainGPU.py:79-96): the q-seat mask from info["seat_idx"],
    # env.step(actions) returning fresh reward / done tensors, terminated |= dones.  +4 B per table-step: the live
    # seat index the step kernel keeps for info["seat_idx"], and 4 B for the reference's prev_invested attribute.
    "poker_hu_dropin": dict(n_players=2, max_players=2, tables=4_194_304, algo_bytes=201, hand_len=16, dropin=True),
}


def peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        return float(json.loads(p.read_text())["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler(threading.Thread):
    """Samples nvidia-smi SM clocks and throttle reasons while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.stop_flag = index, [], threading.Event()

    def run(self):
        try:  # NVML in-process: ~10 ms period instead of one nvidia-smi process per sample
            import pynvml

            pynvml.nvmlInit()
            h = pynvml.nvmlDeviceGetHandleByIndex(self.index)
            mx = pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM)
            R = pynvml
            while not self.stop_flag.is_set():
                sm = pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)
                r = pynvml.nvmlDeviceGetCurrentClocksEventReasons(h) if hasattr(pynvml, "nvmlDeviceGetCurrentClocksEventReasons") \
                    else pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                flags = [bool(r & R.nvmlClocksThrottleReasonHwSlowdown), bool(r & R.nvmlClocksThrottleReasonHwThermalSlowdown),
                         bool(r & R.nvmlClocksThrottleReasonSwThermalSlowdown), bool(r & R.nvmlClocksThrottleReasonSwPowerCap)]
                self.samples.append([str(sm), str(mx)] + ["Active" if f else "Not Active" for f in flags])
                self.stop_flag.wait(0.01)
            return
        except Exception:
            pass
        while not self.stop_flag.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                f = [x.strip() for x in out.strip().split(",")]
                if len(f) >= 6:
                    self.samples.append(f)
            except Exception:
                pass
            self.stop_flag.wait(0.1)

    def sample_once(self):
        """One synchronous sample (taken under load, right before the timed region starts), so the record is never
        empty even when the timed region is shorter than the sampler's period."""
        try:
            import pynvml

            pynvml.nvmlInit()
            h = pynvml.nvmlDeviceGetHandleByIndex(self.index)
            sm = pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)
            mx = pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM)
            r = pynvml.nvmlDeviceGetCurrentClocksEventReasons(h) if hasattr(pynvml, "nvmlDeviceGetCurrentClocksEventReasons") \
                else pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(h)
            flags = [bool(r & pynvml.nvmlClocksThrottleReasonHwSlowdown), bool(r & pynvml.nvmlClocksThrottleReasonHwThermalSlowdown),
                     bool(r & pynvml.nvmlClocksThrottleReasonSwThermalSlowdown), bool(r & pynvml.nvmlClocksThrottleReasonSwPowerCap)]
            self.samples.append([str(sm), str(mx)] + ["Active" if f else "Not Active" for f in flags])
            return
        except Exception:
            pass
        try:
            out = subprocess.run(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                  "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=10).stdout
            f = [x.strip() for x in out.strip().split(",")]
            if len(f) >= 6:
                self.samples.append(f)
        except Exception:
            pass

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        sm = sorted(float(s[0]) for s in self.samples)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(s[2 + i].lower().startswith("active") for s in self.samples)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(self.samples[0][1]), "reasons": reasons,
                "samples": len(sm)}


def workload_config(name, wl, tables):
    """`config` of the JSON line: identical for this repo's arm and for --impl reference (engine details live in
    the line's other keys)."""
    return {"workload": f"{name}: PokerGPU n_players={wl['n_players']}, {tables} tables/GPU, uniform random actions 0..12, "
                        f"whole-batch reset every {wl['hand_len']} steps inside the timed region (reference step "
                        f"convention: every table counted every step)",
            "tables_per_gpu": tables, "hand_len": wl["hand_len"]}


def cpu_baseline(wl, min_seconds=10.0, tables=1 << 20):
    """The oracle port of the reference step (oracle/oracle_poker.c, OpenMP over tables) timed on this
    box's host cores on a bounded sample of the same workload: `tables` tables, whole hands of
    hand_len steps, until min_seconds of stepping have elapsed."""
    import numpy as np
    from oracle import pyoracle
    from pulselib_b200 import handranks

    hr = handranks.load_table()
    avail = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    cores = int(pyoracle.lib().oracle_set_threads(avail))   # torchrun exports OMP_NUM_THREADS=1: set it explicitly
    P, M, L = wl["n_players"], wl["max_players"], wl["hand_len"]
    orc = pyoracle.OraclePoker(hr, n_players=P, max_players=M, n_games=tables, **CFG)
    rng = np.random.default_rng(SEED)
    acts = rng.integers(0, 13, (L, tables)).astype(np.int64)
    steps, spent, hands = 0, 0.0, 0
    while spent < min_seconds:
        orc.tape_decks(SEED, 0, hands)
        t0 = time.perf_counter()
        orc.reset(None)
        orc._sync()
        for i in range(L):
            orc.step_fast(acts[i])
        spent += time.perf_counter() - t0
        steps += L * tables
        hands += 1
    return {"value": steps / spent, "unit": "env-steps/s", "cores": cores, "kind": "port",
            "sample": f"{tables} tables x {hands} hands x {L} steps (reset + step, obs built), oracle/oracle_poker.c "
                      f"with OpenMP on {cores} threads; the reference itself is torch-eager Python and cannot "
                      f"travel to this box"}, spent, steps


def run_reference(args, wl, name):
    """--impl reference: the reference's own CPU path for this step.  The reference is Python
    (torch eager), absent on the GPU box, so the oracle port stands in (kind = "port")."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    per_step = max(1.0, 20.0 / max(1, args.steps + args.warmup))
    tables = int(args.tables or wl["tables"])            # the full configuration, as this repo's arm
    base, spent, steps = cpu_baseline(wl, min_seconds=min(60.0, per_step * (args.steps + args.warmup)), tables=tables)
    line = {"impl": "reference", "metric": "env-steps/sec (PokerGPU NLHE)", "value": base["value"],
            "unit": "env-steps/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * spent / max(1, steps) * tables, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "int32", "data": "synthetic",
            "config": workload_config(name, wl, tables),
            "cpu_baseline": base,
            "e2e": {"value": base["value"], "unit": "env-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def run_env_reference(args, name):
    if int(os.environ.get("RANK", "0")) != 0:
        return
    import bench_envs as be

    base, spent, steps = be.cpu_baseline(name, min_seconds=min(60.0, max(5.0, 2.0 * (args.steps + args.warmup))))
    print(json.dumps({"impl": "reference", "metric": f"env-steps/sec ({name})", "value": base["value"],
                      "unit": "env-steps/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
                      "ms_per_step": None, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                      "dtype": "int32", "data": "synthetic", "config": {"workload": name}, "cpu_baseline": base,
                      "e2e": {"value": base["value"], "unit": "env-steps/s", "h2d_bytes_per_step": 0,
                              "d2h_bytes_per_step": 0}}))


def run_gpu(args, wl, name):
    import torch

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the env-step engine has no CPU fallback")
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    dist = None
    if world > 1:
        import torch.distributed as dist

        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    from pulselib_b200.envs.poker_packed import PokerGPUPacked

    N, L = int(args.tables or wl["tables"]), wl["hand_len"]
    P, M = wl["n_players"], wl["max_players"]
    dropin = bool(wl.get("dropin"))
    if dropin:
        from pulselib_b200 import registration
        from pulselib_b200.gym_compat import gym

        registration.register_all()
        env = gym.make("Pulse-Poker-GPU-v1", device=dev, agents=[], n_players=P, max_players=M, n_games=N, seed=SEED,
                       env_base=rank * N, **CFG)
    else:
        env = PokerGPUPacked(dev, n_players=P, max_players=M, n_games=N, seed=SEED, env_base=rank * N, live_info=False, **CFG)
    gen = torch.Generator(device=dev)
    gen.manual_seed(SEED + rank)
    actions = torch.randint(0, 13, (L, N), dtype=torch.int64, device=dev, generator=gen)
    obs = torch.empty((N, env.obs_size), dtype=torch.float32, device=dev)
    rew = torch.empty(N, dtype=torch.float32, device=dev)
    done = torch.empty(N, dtype=torch.bool, device=dev)
    stream = torch.cuda.current_stream(dev)
    launches = [0]

    loop = {"info": None, "terminated": torch.zeros(N, dtype=torch.bool, device=dev), "done": done}

    def one_step(j):
        if dropin:      # the env-facing part of the reference's loop body
            q_mask = loop["info"]["seat_idx"] == 0
            _, _, d, _, loop["info"] = env.step(actions[j])
            loop["terminated"] |= d
            loop["done"], loop["q_mask"] = d, q_mask
        else:
            env.step_into(actions[j], obs, rew, done)

    def one_reset():
        if dropin:
            _, loop["info"] = env.reset(options={"active_players": False, "rotation": 0})
            loop["terminated"].zero_()
        else:
            env.reset()

    # kernel time: CUDA events around every 5th step launch of the timed region (5 is coprime with the hand length, so
    # every step index of a hand is sampled).  Events around EVERY launch cost 2.9% of the region themselves
    # (profiles/r2ah_event_overhead.txt); PULSE_BENCH_EVENT_EVERY=1 restores that.
    KERNEL_EVENT_EVERY = max(1, int(os.environ.get("PULSE_BENCH_EVENT_EVERY", "5")))

    def run_steps(k, start, pairs=None):
        for i in range(k):
            j = (start + i) % L
            if j == 0:
                one_reset()
                launches[0] += 1
            timed = pairs is not None and i % KERNEL_EVENT_EVERY == 0   # a sample of the launches: two event records per
            if timed:                                                   # step are themselves 2-3% of a 0.12 ms step
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
            one_step(j)
            launches[0] += 1
            if timed:
                e1.record(stream)
                pairs.append((e0, e1))

    def barrier():
        torch.cuda.synchronize(dev)
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize(dev)

    # live-step fraction of the workload (untimed): tables not yet done when each step begins
    one_reset()
    live = 0.0
    for j in range(L):
        live += 1.0 if j == 0 else float((~loop["done"]).float().mean())
        one_step(j)
    live_fraction = live / L

    run_steps(args.warmup, 0)
    sampler = ClockSampler(local)
    sampler.sample_once()        # while the warm-up steps are still running on the device
    barrier()
    sampler.start()
    launches[0] = 0
    pairs = []
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record(stream)
    run_steps(args.steps, args.warmup, pairs)
    t1.record(stream)
    barrier()
    sampler.stop_flag.set()
    sampler.join(timeout=2)
    ms = torch.tensor([t0.elapsed_time(t1)], dtype=torch.float64, device=dev)
    kern_ms = torch.tensor([sum(a.elapsed_time(b) for a, b in pairs) / max(1, len(pairs))], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(kern_ms, op=dist.ReduceOp.MAX)
    total_ms, kernel_ms = float(ms.item()), float(kern_ms.item())
    n_launch = launches[0]

    # End to end with HOST buffers through the repo's host-buffer API (pulselib_b200.host_env.PokerHostEnv ->
    # pulse_poker_host_step): numpy actions in, the reference's float32 observations / rewards / dones out in numpy
    # arrays, every step, host<->device copies inside the timed region.  1 B per table goes in, one packed record per
    # live table comes back and is expanded by native host threads (csrc/host_env.cu).  Host-synchronous API, so the
    # region is timed by wall clock with a device synchronise on both sides.
    import numpy as np

    from pulselib_b200.host_env import PokerHostEnv

    avail = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    local_world = int(os.environ.get("LOCAL_WORLD_SIZE", str(world)))
    threads = args.e2e_threads or max(1, avail // max(1, local_world))
    e2e_steps = min(args.steps, 4 * L)
    # The path's only collective: one more hand on the device with the run statistics accumulated by
    # pulse_stats_accumulate every step, then ONE sum all-reduce of the vector over NCCL, checked against an
    # all_gather of the per-rank vectors.  Outside every timed region.
    from pulselib_b200.sharding import RunStats

    rs = RunStats(dev)
    one_reset()
    for j in range(L):
        one_step(j)
        if dropin:
            pass                                          # the drop-in's step returns fresh reward tensors; not kept here
        else:
            rs.update(rew, done)
    dev_stats = rs.all_reduce()
    if dist is not None and not dropin:
        parts = [torch.zeros_like(rs.vec) for _ in range(world)]
        dist.all_gather(parts, rs.vec)
        want = torch.stack(parts).sum(0)
        # counts are integers held in doubles: exact; the reward sums are doubles added in NCCL's own order (a ring or a
        # tree over the ranks), which from three ranks up is not the order of this sum
        assert torch.equal(want[:2], dev_stats[:2]) and torch.allclose(want[2:], dev_stats[2:], rtol=1e-12, atol=0.0), \
            "NCCL all-reduce of RunStats != sum of gathered vectors"
    if not dropin:
        assert int(dev_stats[0].item()) == world * N * L, "env_steps of the reduced statistics != ranks * tables * steps"
    dev_stats = dev_stats.tolist()
    del obs, rew, done, actions                       # the device-timed leg's buffers
    torch.cuda.empty_cache()
    henv = PokerHostEnv(dev, n_players=P, max_players=M, n_games=N, seed=SEED, env_base=rank * N, chunks=args.e2e_chunks,
                        threads=threads, **CFG)
    h_act = PokerHostEnv.pinned_bytes((L, N))          # pinned host memory, as the contract's inputs are
    h_act[...] = np.random.default_rng(SEED + rank).integers(0, 13, (L, N)).astype(np.uint8)
    h2d = d2h = 0

    def host_steps(k, start, count=False):
        nonlocal h2d, d2h
        for i in range(k):
            j = (start + i) % L
            if j == 0:
                henv.reset()
                if count:
                    d2h += henv.info()["last_d2h_bytes"]
            henv.step(h_act[j])
            if count:
                inf = henv.info()
                h2d += inf["last_h2d_bytes"]; d2h += inf["last_d2h_bytes"]

    host_steps(L, 0)
    barrier()
    henv.timing(reset=True)
    w0 = time.perf_counter()
    host_steps(e2e_steps, 0, count=True)
    torch.cuda.synchronize(dev)
    e2e_ms = torch.tensor([(time.perf_counter() - w0) * 1e3], dtype=torch.float64, device=dev)
    barrier()
    if dist is not None:
        dist.all_reduce(e2e_ms, op=dist.ReduceOp.MAX)
    e2e_value = world * N * e2e_steps / (float(e2e_ms.item()) * 1e-3)
    # the statistics vector of this leg, summed over ranks with the one collective the path has
    stats_local = torch.tensor([float(N * e2e_steps), float(henv.dones.sum()), float(henv.rewards.astype(np.float64).sum())],
                               dtype=torch.float64, device=dev)
    stats_all = stats_local.clone()
    if dist is not None:
        dist.all_reduce(stats_all, op=dist.ReduceOp.SUM)
        gathered = [torch.zeros_like(stats_local) for _ in range(world)]
        dist.all_gather(gathered, stats_local)
        assert torch.allclose(torch.stack(gathered).sum(0), stats_all), "all-reduce of the run statistics disagrees with the gathered sum"
    assert int(stats_all[0].item()) == world * N * e2e_steps
    hinfo = henv.info()
    htime = henv.timing()
    hchunks = [(round(1e3 * a / max(1.0, htime["calls"]), 4), round(1e3 * b / max(1.0, htime["calls"]), 4)) for a, b in henv.chunk_timing()]
    # the same loop with the packed records left unexpanded (pulse_poker_host_step_records): what the link and a
    # record-reading consumer can sustain when the host does not write 64 B of float32 per live table-step
    def record_steps(k):
        for i in range(k):
            if i % L == 0:
                henv.reset(expand=False)
            henv.step_records(h_act[i % L])

    record_steps(L)
    barrier()
    w0 = time.perf_counter()
    record_steps(e2e_steps)
    torch.cuda.synchronize(dev)
    rec_ms = torch.tensor([(time.perf_counter() - w0) * 1e3], dtype=torch.float64, device=dev)
    barrier()
    if dist is not None:
        dist.all_reduce(rec_ms, op=dist.ReduceOp.MAX)
    rec_value = world * N * e2e_steps / (float(rec_ms.item()) * 1e-3)
    henv.close()

    if rank == 0:
        peak, peak_src = peaks()
        A = wl["algo_bytes"]
        achieved = A * N / (kernel_ms * 1e-3) / 1e9
        traffic = None
        tf = ROOT / "profiles" / "traffic.json"
        if tf.exists():
            traffic = json.loads(tf.read_text()).get(name)
        base, _, _ = cpu_baseline(wl, min_seconds=args.cpu_seconds) if world == 1 and args.cpu_seconds > 0 else (None, 0, 0)
        line = {
            "metric": "env-steps/sec (PokerGPU NLHE)", "value": world * N * args.steps / (total_ms * 1e-3),
            "unit": "env-steps/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "int32", "data": "synthetic",
            "config": workload_config(name, wl, N),
            "engine": {"layout": "packed",
                       "api": "gym.make('Pulse-Poker-GPU-v1') + the env-facing loop body of scripts/Poker/trainGPU.py:79-96"
                       if dropin else "PokerGPUPacked.step_into", "parallelism": f"env-sharded x{world}",
                       "l2": "inputs larger than L2 (state+obs+actions per step > 126 MB)"},
            "live_step_fraction": round(live_fraction, 4),
            "live_env_steps_per_s": world * N * args.steps / (total_ms * 1e-3) * live_fraction,
            "e2e": {"value": e2e_value, "unit": "env-steps/s", "h2d_bytes_per_step": h2d / e2e_steps,
                    "d2h_bytes_per_step": d2h / e2e_steps, "steps": e2e_steps, "chunks": hinfo["chunks"],
                    "host_threads": hinfo["threads"], "record_bytes": hinfo["row_bytes"],
                    "api": "pulselib_b200.host_env.PokerHostEnv.step (pulse_poker_host_step): numpy uint8 actions in, "
                           "float32 obs [N,%d] + float32 reward + bool done out in host memory every step, whole-batch "
                           "reset every %d steps; records of live tables only cross PCIe and are expanded by host threads" % (env.obs_size, L),
                    "reference_contract_bytes_per_table_step": 8 + 4 * env.obs_size + 4 + 1,
                    "host_ms_per_step": {k: round(1e3 * v / max(1.0, htime["calls"]), 4) for k, v in htime.items() if k != "calls"},
                    "chunk_ms_seen_done": hchunks,   # per chunk: ms into a call's expansion phase when its records were seen / expanded (last worker)
                    "timing": "wall clock around the host loop (the API is host-synchronous), device synchronised on both sides"},
            "e2e_records": {"value": rec_value, "unit": "env-steps/s",
                            "api": "PokerHostEnv.step_records (pulse_poker_host_step_records): same copies, the packed records stay "
                                   "in pinned host memory, only the done flags are written out"},
            "run_stats_device": dict(zip(("env_steps", "episodes_done", "reward_sum", "reward_sq_sum"), dev_stats)),
            "run_stats_allreduce": {"env_steps": stats_all[0].item(), "done_flags_at_end": stats_all[1].item(),
                                    "reward_sum_at_end": stats_all[2].item(), "ranks": world,
                                    "collective": "nccl all_reduce(sum) of the statistics vector, checked against all_gather" if world > 1 else "single rank"},
            "gpu_launches": n_launch,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "kernel": "poker_packed_step_kernel", "kernel_ms": kernel_ms,
                         "kernel_time": "CUDA events around every %d%s step launch of the timed region (%d samples)" % (KERNEL_EVENT_EVERY, "th" if KERNEL_EVENT_EVERY != 1 else "st", len(pairs)),
                         "algorithmic_bytes_per_table_step": A, "peak_source": peak_src,
                         # the same fraction by the DRAM bytes ncu counted for one launch (profiles/traffic.json): the
                         # packed layout moves fewer bytes than the fixed algorithmic figure, so this is the one that
                         # says how busy HBM really is
                         "frac_by_traffic": (traffic / (kernel_ms * 1e-3) / 1e9 / peak) if traffic else None},
            "clocks": sampler.summary(),
        }
        if base is not None:
            line["cpu_baseline"] = base
        if world == 1 and not args.no_extras:
            line["workloads"] = extra_workloads(dev, rank)
        print(json.dumps(line))
    if dist is not None:
        dist.destroy_process_group()


def run_env_workload(args, name):
    """--workload tfe_4x4 | tetris | particle2d | blackjack: the same JSON line for another env's step kernel."""
    import torch

    import bench_envs as be

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the env-step engine has no CPU fallback")
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    dist = None
    if world > 1:
        import torch.distributed as dist

        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    spec = be.SPECS[name]
    sampler = ClockSampler(local)
    sampler.start()
    m = be.measure(name, dev, rank, world, args.steps, args.warmup, n=args.tables or None, dist=dist)
    sampler.stop_flag.set()
    sampler.join(timeout=2)
    e = be.e2e(name, dev, rank, min(args.steps, 64), n=args.tables or None, chunks=args.e2e_chunks)
    t = torch.tensor([m["total_ms"], m["kernel_ms"], e["ms"]], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms, kernel_ms, e2e_ms = (float(x) for x in t.tolist())
    if rank == 0:
        peak, peak_src = peaks()
        n, A = m["n"], spec["algo_bytes"]
        achieved = A * n / (kernel_ms * 1e-3) / 1e9
        traffic = None
        tf = ROOT / "profiles" / "traffic.json"
        if tf.exists():
            traffic = json.loads(tf.read_text()).get(name)
        line = {
            "metric": f"env-steps/sec ({name})", "value": world * n * args.steps / (total_ms * 1e-3),
            "unit": "env-steps/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32" if name == "particle2d" else "int32", "data": "synthetic",
            "config": {"workload": f"{name}: " + spec["desc"].format(n=n), "envs_per_gpu": n, "launch_mode": m["mode"],
                       "parallelism": f"env-sharded x{world}",
                       "l2": f"{m['replicas']} independent env batch(es) stepped in rotation so each step's inputs were "
                             f"evicted since last touched (working set {m['replicas'] * A * n / 1e6:.0f} MB vs 126 MB L2)"},
            "e2e": {"value": world * e["value"] * (e["ms"] / e2e_ms), "unit": "env-steps/s",
                    "h2d_bytes_per_step": e["h2d_bytes_per_step"], "d2h_bytes_per_step": e["d2h_bytes_per_step"],
                    "steps": e["steps"], "chunks": e["chunks"],
                    **{k: e[k] for k in ("api", "host_threads") if k in e},
                    "timing": "wall clock around the host loop, all streams synchronised"},
            "gpu_launches": m["launches"],
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "frac_by_traffic": (traffic / (kernel_ms * 1e-3) / 1e9 / peak) if traffic else None,
                         "kernel": spec["kernel"], "kernel_ms": kernel_ms,
                         "kernel_time": m["kernel_time"], "algorithmic_bytes_per_env_step": A, "peak_source": peak_src},
            "clocks": sampler.summary(),
        }
        if spec.get("moved_bytes"):      # the fixed algorithmic figure counts bytes this layout never moves: say so
            mb = spec["moved_bytes"]
            line["roofline"]["moved_bytes_per_env_step"] = mb
            line["roofline"]["frac_by_moved_bytes"] = mb * n / (kernel_ms * 1e-3) / 1e9 / peak
        if m.get("live_fraction") is not None:
            line["live_step_fraction"] = round(m["live_fraction"], 4)
            line["live_env_steps_per_s"] = line["value"] * m["live_fraction"]
        if world == 1 and args.cpu_seconds > 0:
            line["cpu_baseline"] = be.cpu_baseline(name, args.cpu_seconds)[0]
        print(json.dumps(line))
    if dist is not None:
        dist.destroy_process_group()


def extra_workloads(dev, rank, steps=256):
    """Compact device-timed numbers of the other env kernels for the default line's `workloads` block."""
    import bench_envs as be

    peak, _ = peaks()
    out = {}
    for name in ("tfe_4x4", "tfe_4x4_rollout", "tetris", "particle2d", "blackjack", "poker_hu_autoreset", "poker_hu_scripted",
                 "poker_hu_rollout", "poker_hu_rollout_tape"):
        try:
            m = be.measure(name, dev, rank, 1, steps, 3)
            A = be.SPECS[name]["algo_bytes"]
            ach = A * m["n"] / (m["kernel_ms"] * 1e-3) / 1e9
            out[name] = {"value": m["value"], "unit": "env-steps/s", "envs": m["n"], "steps": steps,
                         "ms_per_step": m["total_ms"] / steps, "launch_mode": m["mode"],
                         "roofline": {"achieved": ach, "frac": ach / peak, "unit": "GB/s",
                                      "algorithmic_bytes_per_env_step": A, "kernel": be.SPECS[name]["kernel"]}}
            if be.SPECS[name].get("moved_bytes"):
                mb = be.SPECS[name]["moved_bytes"]
                out[name]["roofline"]["moved_bytes_per_env_step"] = mb
                out[name]["roofline"]["frac_by_moved_bytes"] = mb * m["n"] / (m["kernel_ms"] * 1e-3) / 1e9 / peak
            if m.get("live_fraction") is not None:
                out[name]["live_step_fraction"] = round(m["live_fraction"], 4)
                out[name]["live_env_steps_per_s"] = m["value"] * m["live_fraction"]
            if name == "tfe_4x4":      # the metric's second env: end to end and the CPU port beside it
                e = be.e2e(name, dev, rank, 64)
                out[name]["e2e"] = {"value": e["value"], "unit": "env-steps/s", "h2d_bytes_per_step": e["h2d_bytes_per_step"],
                                    "d2h_bytes_per_step": e["d2h_bytes_per_step"], "steps": e["steps"], "chunks": e["chunks"],
                                    **{k: e[k] for k in ("api", "host_threads") if k in e}}
                out[name]["cpu_baseline"] = be.cpu_baseline(name, 5.0)[0]
        except Exception as ex:  # an extra must never take the headline down with it
            out[name] = {"error": f"{type(ex).__name__}: {ex}"}
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=1024)
    ap.add_argument("--warmup", type=int, default=32)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="poker_hu",
                    choices=sorted(WORKLOADS) + ["tfe_4x4", "tfe_4x4_rollout", "tetris", "particle2d", "blackjack", "blackjack_per_step", "poker_hu_autoreset", "poker_hu_scripted", "poker_hu_train_loop", "poker_hu_rollout", "poker_hu_rollout_tape"])
    ap.add_argument("--no-extras", action="store_true", help="skip the `workloads` block of the poker line")
    ap.add_argument("--tables", type=int, default=0, help="tables per GPU (default: the workload's)")
    ap.add_argument("--e2e-chunks", type=int, default=8)
    ap.add_argument("--e2e-threads", type=int, default=0, help="host expander threads per rank (0 = online CPUs / local ranks)")
    ap.add_argument("--cpu-seconds", type=float, default=10.0, help="CPU-baseline sample length (0 = skip)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    if args.workload not in WORKLOADS:
        if args.impl == "reference":
            run_env_reference(args, args.workload)
        else:
            run_env_workload(args, args.workload)
        return
    wl = WORKLOADS[args.workload]
    if args.impl == "reference":
        run_reference(args, wl, args.workload)
    else:
        run_gpu(args, wl, args.workload)


if __name__ == "__main__":
    main()
