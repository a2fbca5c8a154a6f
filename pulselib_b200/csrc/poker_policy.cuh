// Scripted PokerGPU opponents and the build_actions router as per-table rules.
//
// The reference builds the action vector of every step in Python: build_actions
// (/root/reference/environments/Poker/utils.py:108-123) groups seats by agent type, masks the batch per
// type and calls the agent's `action(state[mask])`; the four scripted GPU agents
// (/root/reference/environments/Poker/Player.py:79-176) are rank rules on the acting seat's hole cards
// (obs columns 5..6, `% 13`) and, for SmallBall, the pot (column 9).  Each table's action depends only on
// its own row and its own random draw, so here it is one function of (kind, card0, card1, pot, two chance
// words), evaluated by one thread per table -- standalone on the observation tensor
// (pulse_poker_policy) or inside the packed step kernel on the registers it already holds.
//
// Chance: the reference draws `torch.randint(lo, hi, (n,))` (raise size) and, for LoosePassive,
// `torch.rand(n)`.  Here word 0 of the table's chance pair gives the bounded integer
// lo + mulhi32(w0, hi - lo) and word 1 the uniform float (w1 >> 8) * 2^-24; the pair comes from the tape
// (stream PULSE_STREAM_ACTION, counter = step) or is injected, which is how the reference's agents are
// driven with the same draws for the golden vectors.
#pragma once
#include "common.cuh"

// agent kinds: PULSE_AGENT_* in include/pulse_envs.h
struct PulseSeatKinds { uint8_t kind[16]; };

PULSE_HD int pp_mod13(int v) {  // Python's % on possibly negative ints (cards are 1..52 on every real row)
  int r = v % 13;
  return r < 0 ? r + 13 : r;
}
PULSE_HD long long pp_raise(uint32_t w0, int lo, int hi) {  // raise_distribution = arange(2, 11): index k -> action k + 2
  return 2 + lo + (int)pulse_mulhi32(w0, (uint32_t)(hi - lo));
}

// Action of a scripted seat.  card0/card1: the acting seat's hole cards as the observation carries them;
// pot: observation column 9.  Returns -1 for PULSE_AGENT_EXTERNAL (caller keeps its own action).
PULSE_HD long long pulse_policy_action(int kind, int card0, int card1, float pot, uint32_t w0, uint32_t w1) {
  if (kind == PULSE_AGENT_EXTERNAL) return -1;
  if (kind == PULSE_AGENT_RANDOM) return (long long)pulse_mulhi32(w0, 13u);
  const int r1 = pp_mod13(card0), r2 = pp_mod13(card1);
  const int gap = r1 > r2 ? r1 - r2 : r2 - r1;
  const bool pair = r1 == r2;
  if (kind == PULSE_AGENT_HEURISTIC_HANDS) {                       // Player.py:86-101
    const bool fold = r1 < 8 && r2 < 8;
    const bool high = r1 >= 10 || r2 >= 10;
    return ((pair || high) && !fold) ? pp_raise(w0, 0, 9) : 0;
  }
  if (kind == PULSE_AGENT_TIGHT_AGGRESSIVE) {                      // Player.py:111-124, default 1 (call)
    const bool fold = r1 < 7 && r2 < 7 && gap > 5;
    const bool high = (r1 >= 10 && r2 > 5) || (r2 >= 10 && r1 > 5);
    if ((pair || high) && !fold) return pp_raise(w0, 5, 9);
    return fold ? 0 : 1;
  }
  if (kind == PULSE_AGENT_LOOSE_PASSIVE) {                         // Player.py:134-148, default 0
    const bool fold = r1 <= 4 && r2 <= 4 && gap > 9;
    const bool pairhi = pair && r1 > 8;
    const bool high = (r1 >= 11 && r2 > 9) || (r2 >= 11 && r1 > 9);
    const bool call = (pairhi || high) && !fold;
    const float prob = (float)(w1 >> 8) * 5.9604644775390625e-8f;  // 2^-24
    if (call && prob > 0.9f) return pp_raise(w0, 0, 4);
    return call ? 1 : 0;
  }
  if (kind == PULSE_AGENT_SMALL_BALL) {                            // Player.py:158-174, default 0
    const bool fold = (r1 < 6 && r2 < 6 && pot > 30.0f) || (r1 < 9 && r2 < 9 && pot > 80.0f);
    const bool high = (r1 >= 10 && r2 > 5) || (r2 >= 10 && r1 > 5);
    return ((pair || high) && !fold) ? pp_raise(w0, 0, 3) : 0;
  }
  return -1;
}
