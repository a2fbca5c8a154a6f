// Blackjack: per-env bodies of the batched hit/stand kernels.  Behaviour follows
// /root/reference/environments/blackjack/blackjack.py (reset :23-48, deal_starting_cards :51-101,
// execute_actions :113-162, calculate_rewards :164-177, step :179-186, get_obs :103-108); line cites
// are to that file.  Quirks kept: every ace of the opening deal counts 11 and a two-ace hand loses
// its usable flag (:93-95); the dealer stands on soft 17 (:139, :160); a push is a player win (:174);
// there is no natural-blackjack rule and no auto-reset.
//
// HBM layout (one thread per env):
//   st    uint32 [N]      player_sum:6 | dealer_sum:6 <<6 | deck_pos:6 <<12 | players_card_idx:5 <<18 |
//                         has_ace <<23 | dealer_has_ace <<24 | terminated <<25 | dealer_upcard:4 <<26
//                         (dealer_card_idx = deck_pos - players_card_idx: every drawn card went to one of them)
//   deck  uint8  [52][N]  position-major, so lanes that draw from the same position read adjacent bytes
// The reference's per-card arrays (players_cards, dealer_cards) are a pure function of the deck and
// the two counters; bj_unpack_one replays the draws to rebuild them for inspection and parity tests.
#pragma once
#include "common.cuh"

struct BjState {
  int psum, dsum, pos, pidx, upcard;
  bool ace, dace, term;
};

PULSE_HD uint32_t bj_pack(const BjState& s) {
  return (uint32_t)s.psum | ((uint32_t)s.dsum << 6) | ((uint32_t)s.pos << 12) | ((uint32_t)s.pidx << 18) |
         ((uint32_t)s.ace << 23) | ((uint32_t)s.dace << 24) | ((uint32_t)s.term << 25) | ((uint32_t)s.upcard << 26);
}
PULSE_HD BjState bj_unpack_word(uint32_t w) {
  BjState s;
  s.psum = w & 63; s.dsum = (w >> 6) & 63; s.pos = (w >> 12) & 63; s.pidx = (w >> 18) & 31;
  s.ace = (w >> 23) & 1; s.dace = (w >> 24) & 1; s.term = (w >> 25) & 1; s.upcard = (w >> 26) & 15;
  return s;
}

PULSE_HD int bj_value(int card) {  // clamp(card % 13 + 1, max=10)  :53
  int v = card % 13 + 1;
  return v > 10 ? 10 : v;
}

// one draw into (sum, usable-ace flag): ace counts 11 only when no usable ace is held (:119-135, :143-157)
PULSE_HD int bj_draw(int card, int& sum, bool& ace) {
  int r = bj_value(card);
  const bool first_ace = r == 1 && !ace;
  if (first_ace) r = 11;
  ace = ace || first_ace;
  sum += r;
  if (sum > 21 && ace) { sum -= 10; ace = false; }
  return r;
}

// Fisher-Yates deck of 0..51 from the chance tape (same construction as the poker deck)
PULSE_HD void bj_tape_deck(uint8_t (&d)[52], uint64_t seed, uint64_t env_id, uint32_t episode) {
  for (int i = 0; i < 52; ++i) d[i] = (uint8_t)i;
  PulseU4 blk; blk.x = blk.y = blk.z = blk.w = 0;
  for (int i = 0; i < 52; ++i) {
    if ((i & 3) == 0) blk = pulse_tape(seed, env_id, PULSE_STREAM_BJ_DECK, episode * 16u + (uint32_t)(i >> 2));
    const uint32_t u = (i & 3) == 0 ? blk.x : (i & 3) == 1 ? blk.y : (i & 3) == 2 ? blk.z : blk.w;
    const int j = i + (int)pulse_mulhi32(u, (uint32_t)(52 - i));
    const uint8_t t = d[i]; d[i] = d[j]; d[j] = t;
  }
}

// opening deal (:51-101): player, dealer (up-card), player, dealer from deck positions 0..3
PULSE_HD BjState bj_deal(int c0, int c1, int c2, int c3) {
  int p1 = bj_value(c0), d1 = bj_value(c1), p2 = bj_value(c2), d2 = bj_value(c3);
  const bool pa1 = p1 == 1, da1 = d1 == 1, pa2 = p2 == 1, da2 = d2 == 1;
  if (pa1) p1 = 11;
  if (da1) d1 = 11;
  if (pa2) p2 = 11;
  if (da2) d2 = 11;
  BjState s;
  s.ace = pa1 || pa2; s.dace = da1 || da2;            // :58, :78, :69, :82-85
  s.psum = p1 + p2; s.dsum = d1 + d2;                 // :89-90
  if (s.psum > 21 && s.ace) { s.psum -= 10; s.ace = false; }    // :93-95
  if (s.dsum > 21 && s.dace) { s.dsum -= 10; s.dace = false; }  // :99-101
  s.pos = 4; s.pidx = 2; s.upcard = d1; s.term = false;
  return s;
}

// reset of one env: decks come from `injected` (int32/int64-free: uint8 [N,52] row-major) or the tape
PULSE_HD BjState bj_reset_state(uint32_t* __restrict__ st, uint8_t* __restrict__ deck, const uint8_t* __restrict__ injected,
                                uint64_t seed, uint64_t env_base, uint32_t episode, long long g, long long n) {
  uint8_t d[52];
  if (injected) { for (int i = 0; i < 52; ++i) d[i] = injected[g * 52 + i]; }
  else bj_tape_deck(d, seed, env_base + (uint64_t)g, episode);
  for (int i = 0; i < 52; ++i) deck[(long long)i * n + g] = d[i];
  const BjState s = bj_deal(d[0], d[1], d[2], d[3]);
  st[g] = bj_pack(s);
  return s;
}
PULSE_HD void bj_reset_one(uint32_t* __restrict__ st, uint8_t* __restrict__ deck, const uint8_t* __restrict__ injected,
                           uint64_t seed, uint64_t env_base, uint32_t episode, int32_t* __restrict__ obs,
                           int32_t* __restrict__ rewards, uint8_t* __restrict__ terminated, long long g, long long n) {
  const BjState s = bj_reset_state(st, deck, injected, seed, env_base, episode, g, n);
  obs[3 * g] = s.psum; obs[3 * g + 1] = s.ace; obs[3 * g + 2] = s.upcard;
  rewards[g] = 0; terminated[g] = 0;
}

// one step on a state held in registers (:179-186); returns the reward
PULSE_HD int bj_step_state(BjState& s, long long a, const uint8_t* __restrict__ deck, long long g, long long n) {
  int reward = 0;                                      // :183 rewards.zero_()
  if (!s.term) {
    if (a == 0) {                                      // hit :117-135
      bj_draw(deck[(long long)s.pos * n + g], s.psum, s.ace);
      s.pos += 1; s.pidx += 1;
      if (s.psum > 21) { reward = -1; s.term = true; } // :166-168
    } else if (a == 1) {                               // stand :138-160
      while (s.dsum < 17) {                            // (after a draw the <= 21 test is implied by < 17)
        bj_draw(deck[(long long)s.pos * n + g], s.dsum, s.dace);
        s.pos += 1;
      }
      reward = (s.dsum > 21 || s.psum >= s.dsum) ? 1 : -1;  // :171-176 push counts as a win
      s.term = true;                                   // :177
    }
  }
  return reward;
}

// step of one env (:179-186)
PULSE_HD void bj_step_one(uint32_t* __restrict__ st, const uint8_t* __restrict__ deck, const int64_t* __restrict__ actions,
                          int32_t* __restrict__ obs, int32_t* __restrict__ rewards, uint8_t* __restrict__ terminated,
                          long long g, long long n) {
  const uint32_t w0 = st[g];
  BjState s = bj_unpack_word(w0);
  const int reward = bj_step_state(s, actions[g], deck, g, n);
  const uint32_t w1 = bj_pack(s);
  if (w1 != w0) st[g] = w1;
  obs[3 * g] = s.psum; obs[3 * g + 1] = s.ace; obs[3 * g + 2] = s.upcard;  // :103-108
  rewards[g] = reward;
  terminated[g] = s.term;
}

// Rebuilds the reference-shaped tensors of one env from (st, deck).  All outputs int32 except
// decks (int64, as torch.argsort produces) and the three uint8 flags.
struct BjCompat {
  int64_t* decks;             // [N,52]
  int32_t* deck_positions;    // [N]
  int32_t* players_cards;     // [N,20]
  int32_t* players_card_idx;  // [N]
  int32_t* player_card_sums;  // [N]
  uint8_t* has_ace;           // [N]
  int32_t* dealer_cards;      // [N,20]
  int32_t* dealer_card_idx;   // [N]
  int32_t* dealer_upcard;     // [N]
  int32_t* dealer_card_sums;  // [N]
  uint8_t* dealer_has_ace;    // [N]
};

PULSE_HD void bj_unpack_one(const uint32_t* __restrict__ st, const uint8_t* __restrict__ deck, const BjCompat& C,
                            long long g, long long n) {
  const BjState s = bj_unpack_word(st[g]);
  for (int i = 0; i < 52; ++i) C.decks[g * 52 + i] = deck[(long long)i * n + g];
  int32_t* pc = C.players_cards + g * 20;
  int32_t* dc = C.dealer_cards + g * 20;
  for (int i = 0; i < 20; ++i) pc[i] = dc[i] = 0;
#define BJ_CARD(p) ((int)deck[(long long)(p) * n + g])
  // opening cards are stored with every ace as 11 (:53-85)
  const int o[4] = {bj_value(BJ_CARD(0)), bj_value(BJ_CARD(1)), bj_value(BJ_CARD(2)), bj_value(BJ_CARD(3))};
  pc[0] = o[0] == 1 ? 11 : o[0]; dc[0] = o[1] == 1 ? 11 : o[1];
  pc[1] = o[2] == 1 ? 11 : o[2]; dc[1] = o[3] == 1 ? 11 : o[3];
  BjState r = bj_deal(BJ_CARD(0), BJ_CARD(1), BJ_CARD(2), BJ_CARD(3));
  int pos = 4;
  for (int k = 2; k < s.pidx && k < 20; ++k) pc[k] = bj_draw(BJ_CARD(pos++), r.psum, r.ace);   // hits come first
  const int didx = s.pos - s.pidx;
  for (int k = 2; k < didx && k < 20; ++k) dc[k] = bj_draw(BJ_CARD(pos++), r.dsum, r.dace);    // then the dealer
#undef BJ_CARD
  C.deck_positions[g] = s.pos; C.players_card_idx[g] = s.pidx; C.player_card_sums[g] = s.psum;
  C.has_ace[g] = s.ace; C.dealer_card_idx[g] = didx; C.dealer_upcard[g] = s.upcard;
  C.dealer_card_sums[g] = s.dsum; C.dealer_has_ace[g] = s.dace;
}
