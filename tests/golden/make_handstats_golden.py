"""Generates tests/golden/poker_handstats.npz with the reference's OWN metric code
(/root/reference/utils/performance.py, imported in place; build container only).

A synthetic but realistic trace of a performance run is built -- per step: done flags, the Q seat's stacks,
stages, buttons -- and pushed through exactly what /root/reference/scripts/Poker/trainGPU_performance.py:186-206
does on the host (terminated_before / newly_done masks, lists of gathered tensors).  The reference's functions
then produce the numbers the device accumulator has to reproduce: summarize_episode_performance_metrics,
calculate_lcb95_bb_per_100_from_tensor, calculate_street_win_percentages, calculate_position_win_rates,
calculate_seat_balanced_bb_per_100, calculate_q_seat_positions, bucketize_terminal_stages.

    python tests/golden/make_handstats_golden.py
"""
import importlib.util
import sys
from pathlib import Path

import numpy as np
import torch

OUT = Path(__file__).resolve().parent / "poker_handstats.npz"
spec = importlib.util.spec_from_file_location("_ref_performance", "/root/reference/utils/performance.py")
perf = importlib.util.module_from_spec(spec)
sys.modules["_ref_performance"] = perf
spec.loader.exec_module(perf)


def main():
    rng = np.random.default_rng(20260401)
    rec = {}
    for tag, N, P, AP, q_seat, hands, steps in (("six", 5000, 6, 5, 2, 3, 14), ("hu", 8000, 2, 2, 0, 2, 9)):
        deltas, stages_l, pos_l = [], [], []
        trace = {k: [] for k in ("done", "stacks", "stages", "button", "initial")}
        for h in range(hands):
            button = rng.integers(0, AP, N).astype(np.int32)
            stacks = rng.integers(20, 200, (N, P)).astype(np.int32)
            initial = stacks[:, q_seat].copy()
            q_pos = perf.calculate_q_seat_positions(torch.from_numpy(button), q_seat=q_seat, active_players=AP)
            terminated = torch.zeros(N, dtype=torch.bool)
            stages = np.zeros(N, np.int32)
            for t in range(steps):
                live = ~terminated.numpy()
                stages = np.where(live, np.minimum(stages + (rng.random(N) < 0.4), 3), stages).astype(np.int32)
                stacks = np.where(live[:, None], stacks + rng.integers(-6, 7, (N, P)), stacks).astype(np.int32)
                dones = torch.from_numpy((rng.random(N) < 0.25) | ~live)      # done tables stay done
                fin = dones.numpy() & live
                stages = np.where(fin & (rng.random(N) < 0.5), rng.integers(4, 6, N), stages).astype(np.int32)
                # --- trainGPU_performance.py:186-206 ---
                terminated_before = terminated.clone()
                terminated |= dones
                newly_done = dones & ~terminated_before
                if newly_done.any():
                    final_q = torch.from_numpy(stacks)[newly_done, q_seat].to(torch.float32)
                    completed = final_q - torch.from_numpy(initial)[newly_done].to(torch.float32)
                    deltas.append(completed)
                    stages_l.append(torch.from_numpy(stages)[newly_done].to(torch.int64).clone())
                    pos_l.append(q_pos[newly_done].clone())
                for k, v in (("done", dones.numpy().astype(np.uint8)), ("stacks", stacks.copy()), ("stages", stages.copy()),
                             ("button", button.copy()), ("initial", initial.copy())):
                    trace[k].append(v)
        summ = perf.summarize_episode_performance_metrics(episode_reward=torch.zeros(()), cumulative_reward=torch.zeros(()),
                                                          hand_bb_deltas=deltas)
        flat = perf.flatten_metric_batches(deltas, dtype=torch.float32, device=torch.device("cpu"))
        street = perf.calculate_street_win_percentages(hand_bb_deltas=deltas, hand_terminal_stages=stages_l)
        posw = perf.calculate_position_win_rates(hand_bb_deltas=deltas, hand_positions=pos_l)
        for k, v in trace.items():
            rec[f"{tag}/{k}"] = np.stack(v)
        rec[f"{tag}/meta"] = np.array([N, P, AP, q_seat, hands, steps], np.int32)
        rec[f"{tag}/summary"] = np.array([float(summ["mean_bb_delta"]), float(summ["hand_win_rate"]),
                                          float(summ["hands_completed"]), float(summ["field_bb_per_100"]),
                                          float(perf.calculate_lcb95_bb_per_100_from_tensor(flat)),
                                          float(perf.calculate_seat_balanced_bb_per_100(hand_bb_deltas=deltas,
                                                                                        hand_positions=pos_l))], np.float64)
        rec[f"{tag}/street_win_pct"] = np.array([float(street[n]) for n in ("preflop", "flop", "turn", "river", "showdown")])
        rec[f"{tag}/position_hands"] = np.array([int(posw.get(f"position_{p}", {}).get("hands", 0)) for p in range(16)])
        rec[f"{tag}/position_wins"] = np.array([int(posw.get(f"position_{p}", {}).get("wins", 0)) for p in range(16)])
        print(tag, rec[f"{tag}/summary"], rec[f"{tag}/street_win_pct"], rec[f"{tag}/position_hands"][:AP])
    np.savez_compressed(OUT, **rec)
    print("wrote", OUT, OUT.stat().st_size)


if __name__ == "__main__":
    main()
