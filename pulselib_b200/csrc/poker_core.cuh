// PokerGPU per-table step logic, shared by the reference-shaped ("compat") kernels and the packed
// throughput kernels.  One thread owns one table; the whole table lives in registers between one
// load and one store.  Semantics follow /root/reference/environments/Poker/PokerGPU.py:527-633
// (SURVEY.md Appendix A); line cites are to that file.
#pragma once
#include "common.cuh"
#include "poker_eval.cuh"
#include <limits.h>

enum { PK_ACTIVE = 0, PK_FOLDED = 1, PK_ALLIN = 2, PK_SITOUT = 3 };  // :11

// phases of a step; the fused kernel runs PK_PH_STEP, the sub-step entry points run one bit
enum {
  PK_PH_EQUITY = 1, PK_PH_EXECUTE = 2, PK_PH_ADVANCE = 4, PK_PH_FOLDWIN = 8, PK_PH_SHOWDOWN = 16,
  PK_PH_CLEAR = 32, PK_PH_REWARD = 64, PK_PH_OBS = 128, PK_PH_STEP = 255,
  PK_PH_NEWLY_IS_DONE = 256,  // resolve_* called on their own: is_done is the "newly done" mask (:620)
  PK_PH_RAW_REWARD = 512      // poker_reward_gpu on its own: no legality masking, actor given
};

struct PokerParams {
  float w1, w2;
  int K, alpha;
  int P, AP, M;  // n_players, active_players, max_players
};

template <int MAXP>
struct PokerTable {
  int32_t stack[MAXP], crb[MAXP], inv[MAXP];
  int32_t hole0[MAXP], hole1[MAXP];
  uint32_t st;  // status bit-planes: bit s = low bit of seat s, bit 16+s = high bit
  int32_t pot, highest, last_raise;
  int32_t board[5];  // -1 = not dealt
  int32_t stage, idx, agg, acted, button, deck_pos;
  int32_t eq_stage;  // packed layout only: the street (1..3) the reference's equity cache was last computed on, 0 = none
  bool done, dirty;
};

template <int MAXP>
PULSE_HD int32_t pk_sel(const int32_t (&a)[MAXP], int i) {
  int32_t v = a[0];
#pragma unroll
  for (int s = 1; s < MAXP; ++s) if (i == s) v = a[s];
  return v;
}
template <int MAXP>
PULSE_HD void pk_put(int32_t (&a)[MAXP], int i, int32_t v) {
#pragma unroll
  for (int s = 0; s < MAXP; ++s) if (i == s) a[s] = v;
}

PULSE_HD int pk_status(uint32_t st, int s) { return ((st >> s) & 1u) | (((st >> (16 + s)) & 1u) << 1); }
PULSE_HD uint32_t pk_set_status(uint32_t st, int s, int v) {
  uint32_t lo = 1u << s, hi = 1u << (16 + s);
  st &= ~(lo | hi);
  if (v & 1) st |= lo;
  if (v & 2) st |= hi;
  return st;
}
PULSE_HD uint32_t pk_seatmask(int n) { return (1u << n) - 1u; }
PULSE_HD uint32_t pk_active_mask(uint32_t st, int n) { return ~(st | (st >> 16)) & pk_seatmask(n); }
PULSE_HD uint32_t pk_allin_mask(uint32_t st, int n) { return (st >> 16) & ~st & pk_seatmask(n); }
PULSE_HD uint32_t pk_live_mask(uint32_t st, int n) { return ~st & pk_seatmask(n); }  // ACTIVE|ALLIN: low bit clear

// first ACTIVE seat scanning (from+1 .. from+AP) % AP, or -1 (:554-561, :215-228)
PULSE_HD int pk_next_active(uint32_t st, int from, int AP) {
  uint32_t a = pk_active_mask(st, AP);
  if (!a) return -1;
  uint32_t above = a & ~((2u << from) - 1u);
  return PULSE_FFS(above ? above : a) - 1;
}

// The reference's own chained walk p = HR[p + card] from 53 (:437-444, :477-478), with torch's wrap of
// negative indices.  Needed only where a seat's cards are NOT distinct valid cards (the reference's
// white-box tests put -1 boards and colliding cards into the state): there the table's answer is
// arbitrary but deterministic, and only the walk itself reproduces it.  Valid hands never come here.
PULSE_HD int32_t pk_hr_at(const int32_t* __restrict__ hr, long long n, long long i) {
  if (i < 0) i += n;
  return (i >= 0 && i < n) ? PULSE_LDG(hr + i) : 0;
}
PULSE_HD int32_t pk_table_walk(const int32_t* __restrict__ hr, long long n, int h0, int h1,
                               const int32_t (&board)[5], int n_board) {
  int32_t p = pk_hr_at(hr, n, 53ll + h0);
  p = pk_hr_at(hr, n, (long long)p + h1);
  for (int i = 0; i < n_board; ++i) p = pk_hr_at(hr, n, (long long)p + board[i]);
  return p;
}
// The whole 52-card deck of (seed, env, hand) on the chance tape: forward Fisher-Yates, position i fixed at
// iteration i (replaces rand().argsort() + 1, PokerGPU.py:86).  The packed deal runs the same loop for the first
// 2 AP + 8 positions only, so its cards are this deck's prefix.
PULSE_HD void pk_tape_deck(uint8_t (&d)[52], uint64_t seed, uint64_t env_id, uint32_t hand) {
  for (int i = 0; i < 52; ++i) d[i] = (uint8_t)(i + 1);
  PulseU4 blk;
  blk.x = blk.y = blk.z = blk.w = 0;
  for (int i = 0; i < 51; ++i) {
    if ((i & 3) == 0) blk = pulse_tape(seed, env_id, PULSE_STREAM_POKER_DECK, hand * 16u + (uint32_t)(i >> 2));
    uint32_t u = (i & 3) == 0 ? blk.x : (i & 3) == 1 ? blk.y : (i & 3) == 2 ? blk.z : blk.w;
    int j = i + (int)pulse_mulhi32(u, (uint32_t)(52 - i));
    uint8_t tmp = d[i]; d[i] = d[j]; d[j] = tmp;
  }
}

PULSE_HD int pk_popc64(uint64_t m) { return PULSE_POPC((uint32_t)m) + PULSE_POPC((uint32_t)(m >> 32)); }

// The reference's hand-strength proxy for one seat (:455-525).  stage 3: 7-card value; stage 2:
// 6-card value; stage 1: HR[HR[p5]] where HR[p5] is the 5-card value, i.e. one real lookup into the
// low part of the table (only its first 36,875 entries are ever touched).
PULSE_HD float pk_equity(int stage, int h0, int h1, const int32_t (&board)[5],
                                           const int32_t* __restrict__ hand_ranks, long long n_hr) {
  if (stage < 1 || stage > 3) return 0.5f;  // :461
  uint64_t m = pe_add_card64(h1, pe_add_card64(h0, 0));
  m = pe_add_card64(board[2], pe_add_card64(board[1], pe_add_card64(board[0], m)));
  if (stage >= 2) m = pe_add_card64(board[3], m);
  if (stage >= 3) m = pe_add_card64(board[4], m);
  int v;
  float lo = 4109.0f, den = 32765.0f;  // :17-18, :65
  if (stage == 1) { lo = 74359.0f; den = 749420.0f; }      // :15-16, :66
  if (pk_popc64(m) == stage + 4) {
    v = pe_eval_mask64(m);
    if (stage == 1) v = pk_hr_at(hand_ranks, n_hr, v);      // :521 (non-negative here)
  } else {  // invalid or repeated cards: the literal walk (:477-478, :497-500, :518-521)
    v = pk_table_walk(hand_ranks, n_hr, h0, h1, board, stage + 2);
    if (stage <= 2) v = pk_hr_at(hand_ranks, n_hr, v);
    if (stage == 1) v = pk_hr_at(hand_ranks, n_hr, v);
  }
  float e = PULSE_FDIV(PULSE_FSUB((float)v, lo), den);
  return fminf(fmaxf(e, 0.0f), 1.0f);
}

// execute_actions for the seat to act (:230-303).  Written without branches on the action kind: fold, call and the
// raises of one warp's tables otherwise run as three instruction streams one after another (13-17 of 32 lanes per
// instruction in the r2b profile); every table computes the raise amount and the bet, selects decide what is kept.
template <int MAXP>
PULSE_HD void pk_execute(PokerTable<MAXP>& t, long long a) {
  const int idx = t.idx;
  // acts only if ACTIVE and the table is not done (:233); negative ids match no mask in the reference
  const bool act = pk_status(t.st, idx) == PK_ACTIVE && !t.done && a >= 0;
  if (!act) return;
  int stack = pk_sel(t.stack, idx), crb = pk_sel(t.crb, idx), inv = pk_sel(t.inv, idx);
  const int call_cost = t.highest - crb;    // :232
  const bool fold = a == 0;                 // :236-238
  const bool raise = a >= 2;                // :255-303 (ids above 12 raise by nothing)
  float f = 0.25f;                          // raise_fractions (:39), ids 3..11
  f = a == 4 ? 0.33f : f; f = a == 5 ? 0.50f : f; f = a == 6 ? 0.75f : f; f = a == 7 ? 1.00f : f;
  f = a == 8 ? 1.50f : f; f = a == 9 ? 2.00f : f; f = a == 10 ? 3.00f : f; f = a == 11 ? 4.00f : f;
  int raise_amt = PULSE_F2I_RZ(PULSE_FMUL((float)t.pot, f));   // :270 one fp32 multiply, truncation
  raise_amt = a == 2 ? t.last_raise : raise_amt;                // :259
  raise_amt = a == 12 ? stack : raise_amt;                      // :264
  raise_amt = (a < 2 || a > 12) ? 0 : raise_amt;
  const int want = call_cost + raise_amt;                        // a == 1: call_cost (:241-252); :272-273
  int bet = want < stack ? want : stack;
  bet = fold ? 0 : bet;
  const bool is_raise = raise && bet > call_cost;                // :274-275
  stack -= bet; crb += bet; inv += bet; t.pot += bet;
  if (fold) t.st = pk_set_status(t.st, idx, PK_FOLDED);
  else if (stack == 0) t.st = pk_set_status(t.st, idx, PK_ALLIN);
  pk_put(t.stack, idx, stack); pk_put(t.crb, idx, crb); pk_put(t.inv, idx, inv);
  if (is_raise) {  // :288-301
    const int size = crb - t.highest;
    t.highest = crb;
    if (size >= t.last_raise) { t.agg = idx; t.acted = 0; t.last_raise = size; }
  }
  t.acted += 1;  // :238, :252, :303
}

// next actor, round close, street transition (:547-616).  Deck::card(pos) returns the card at a
// deck position of this table.
template <int MAXP, class Deck>
PULSE_HD void pk_advance(PokerTable<MAXP>& t, const PokerParams& pr, const Deck& deck) {
  const int P = pr.P, AP = pr.AP;
  int truly_active = PULSE_POPC(pk_active_mask(t.st, P));      // :547
  bool all_acted = t.acted >= truly_active;                // :549
  bool round_over = t.done || truly_active == 0;           // :552-553
  int nxt = pk_next_active(t.st, t.idx, AP);               // :554-561
  if (!round_over)                                         // :562-569
    round_over = nxt < 0 || (all_acted && t.idx == t.agg) || (all_acted && nxt == t.agg);
  if (!round_over && nxt >= 0) t.idx = nxt;                // :571-573
  int lc = PULSE_POPC(pk_live_mask(t.st, P));                  // :576
  bool early = lc <= 1 && round_over;                      // :577
  if (early) t.done = true;                                // :578
  if (round_over && !early && !t.done) {                   // :580
    t.last_raise = 1;                                      // :581
    t.stage += 1;                                          // :584
    t.highest = 0;                                         // :585
    {  // (button + 1) % AP (:586) without the division: button is a seat in [0, AP) on every path that deals
      int b1 = t.button + 1;
      if (b1 >= AP) { b1 -= AP; if (b1 >= AP) b1 %= AP; }  // second branch: hand-written states only
      t.agg = b1;
    }
    t.acted = 0;                                           // :587
#pragma unroll
    for (int s = 0; s < MAXP; ++s) t.crb[s] = 0;           // :588
    int first = pk_next_active(t.st, t.button, AP);        // :589
    if (first >= 0) t.idx = first;
    if (t.stage > 3) { t.done = true; t.stage = 4; }       // :595-598
    else if (t.stage == 1) {                               // :601-604
      t.deck_pos += 1;
      t.board[0] = deck.card(t.deck_pos); t.board[1] = deck.card(t.deck_pos + 1); t.board[2] = deck.card(t.deck_pos + 2);
      t.deck_pos += 3; t.dirty = true;
    } else if (t.stage == 2) {                             // :607-610
      t.deck_pos += 1; t.board[3] = deck.card(t.deck_pos); t.deck_pos += 1; t.dirty = true;
    } else {                                               // :613-616
      t.deck_pos += 1; t.board[4] = deck.card(t.deck_pos); t.deck_pos += 1; t.dirty = true;
    }
  }
}

// resolve_fold_winners (:331-338): exactly one live seat takes the pot
template <int MAXP>
PULSE_HD void pk_fold_winner(PokerTable<MAXP>& t, const PokerParams& pr) {
  uint32_t live = pk_live_mask(t.st, pr.P);
  if (PULSE_POPC(live) != 1) return;
  int w = PULSE_FFS(live) - 1;
  pk_put(t.stack, w, pk_sel(t.stack, w) + t.pot);
  t.pot = 0;
}

// resolve_terminated_games + _award_showdown_side_pots (:380-453, :340-378)
// Ranker::rank(t, s) gives the showdown value of live seat s once the board is complete.
// Reference-shaped layout: computed evaluator for seven distinct valid cards, the literal table walk otherwise.
struct PokerFaithfulRanker {
  const int32_t* hr; long long n;
  template <int MAXP>
  PULSE_HD int32_t rank(const PokerTable<MAXP>& t, int s) const {
    const int h0 = pk_sel(t.hole0, s), h1 = pk_sel(t.hole1, s);
    uint64_t m = pe_add_card64(h1, pe_add_card64(h0, 0));
#pragma unroll
    for (int i = 0; i < 5; ++i) m = pe_add_card64(t.board[i], m);
    if (pk_popc64(m) == 7) return pe_eval_mask64(m);
    return pk_table_walk(hr, n, h0, h1, t.board, 5);
  }
};

struct PokerCardRanker {  // evaluates the seat's two hole cards + the five board cards
  template <int MAXP>
  PULSE_HD int32_t rank(const PokerTable<MAXP>& t, int s) const {
    uint64_t m = 0;
#pragma unroll
    for (int i = 0; i < 5; ++i) m = pe_add_card64(t.board[i], m);
    m = pe_add_card64(pk_sel(t.hole0, s), m);
    m = pe_add_card64(pk_sel(t.hole1, s), m);
    return pe_eval_mask64(m);
  }
};

template <int MAXP, class Deck, class Ranker>
PULSE_HD void pk_showdown(PokerTable<MAXP>& t, const PokerParams& pr, const Deck& deck, const Ranker& ranker) {
  const int P = pr.P, AP = pr.AP;
  if (!(t.stage < 5)) return;                               // :386
  uint32_t live = pk_live_mask(t.st, P);
  if (!(PULSE_POPC(live) > 1)) return;                          // :390
  if (t.stage == 0) {                                       // :397-405
    t.deck_pos += 1;
    t.board[0] = deck.card(t.deck_pos); t.board[1] = deck.card(t.deck_pos + 1); t.board[2] = deck.card(t.deck_pos + 2);
    t.deck_pos += 3;
  }
  if (t.stage <= 1) { t.deck_pos += 1; t.board[3] = deck.card(t.deck_pos); t.deck_pos += 1; }  // :407-411
  if (t.stage <= 2) { t.deck_pos += 1; t.board[4] = deck.card(t.deck_pos); t.deck_pos += 1; }  // :416-419
  if (MAXP == 2) {
    // Two seats, both live: the side-pot layers of :346-378 in closed form.  Layer 1 (up to the smaller investment,
    // two contributors) goes to the better hand, split evenly on a tie (2 lo is even: no remainder); layer 2 (the
    // excess of the bigger investment, one contributor, who is live) goes back to that seat.
    const int32_t r0 = ranker.rank(t, 0), r1 = ranker.rank(t, 1);
    const int32_t i0 = t.inv[0], i1 = t.inv[1];
    const int32_t lo = i0 < i1 ? i0 : i1, hi = i0 < i1 ? i1 : i0;
    if (lo > 0) {
      t.stack[0] += r0 > r1 ? 2 * lo : r0 == r1 ? lo : 0;
      t.stack[1] += r1 > r0 ? 2 * lo : r0 == r1 ? lo : 0;
    }
    if (i0 > i1) t.stack[0] += hi - lo;
    if (i1 > i0) t.stack[1] += hi - lo;
    t.pot = 0; t.stage = 5;                                 // :452-453
    return;
  }
  int32_t rank[MAXP];
#pragma unroll
  for (int s = 0; s < MAXP; ++s) {                          // :427-450
    rank[s] = INT_MIN;
    if (s < AP && ((live >> s) & 1u)) rank[s] = ranker.rank(t, s);
  }
  // side pots: one layer per distinct investment level, ascending (:346-378).  Iterating seats as
  // "level = inv[k]" and skipping levels already handled equals the reference's sorted layers:
  // duplicate levels have size 0 there.
  int32_t prev_level = 0;
  for (int layer = 0; layer < AP; ++layer) {
    int32_t level = INT_MAX;  // smallest investment above prev_level
#pragma unroll
    for (int s = 0; s < MAXP; ++s)
      if (s < AP && t.inv[s] > prev_level && t.inv[s] < level) level = t.inv[s];
    if (level == INT_MAX) break;
    int contributors = 0, winners = 0, first = -1;
    int32_t best = INT_MIN;
#pragma unroll
    for (int s = 0; s < MAXP; ++s)
      if (s < AP && t.inv[s] >= level) {
        contributors++;
        if (((live >> s) & 1u) && rank[s] > best) best = rank[s];
      }
#pragma unroll
    for (int s = MAXP - 1; s >= 0; --s)
      if (s < AP && t.inv[s] >= level && ((live >> s) & 1u) && rank[s] == best) { winners++; first = s; }
    int32_t pot = (level - prev_level) * contributors;      // :359-360
    if (winners > 0) {                                      // :362
      int32_t share = winners == 1 ? pot : winners == 2 ? (pot >> 1) : pot / winners;  // pot >= 0
      int32_t rem = pot - share * winners;
#pragma unroll
      for (int s = 0; s < MAXP; ++s)
        if (s < AP && t.inv[s] >= level && ((live >> s) & 1u) && rank[s] == best)
          t.stack[s] += share + (s == first ? rem : 0);     // :374-378
    }
    prev_level = level;
  }
  t.pot = 0; t.stage = 5;                                   // :452-453
}

// poker_reward_gpu (:305-329), fp32 with one rounding per op like the eager reference
template <int MAXP>
PULSE_HD float pk_reward(const PokerTable<MAXP>& t, const PokerParams& pr, long long a,
                                           float e, int prev_inv) {
  const int cnt = PULSE_POPC(pk_live_mask(t.st, pr.P));           // :309
  // :310 fair = 1 / max(cnt, 1): for one or two live seats the quotient is exact, no division needed
  const float fair = cnt <= 1 ? 1.0f : cnt == 2 ? 0.5f : PULSE_FDIV(1.0f, (float)cnt);
  int cc = t.highest - prev_inv; if (cc < 0) cc = 0;                     // :311
  float pot = (float)t.pot;
  float m = PULSE_FMUL(e, pot);                               // :315
  // :316; nothing to call is the common case and 0 / x is exactly 0 -- and a zero numerator sends the IEEE division
  // down its slow-path subroutine (5% of the step's instructions in the r2b profile)
  float o = cc == 0 ? 0.0f : PULSE_FDIV((float)cc, PULSE_FADD((float)(t.pot + cc), 1e-6f));
  float s = 0.0f;
  if (a == 1) s = PULSE_FMUL(PULSE_FSUB(e, o), pot);           // :326
  else if (a == 0) s = PULSE_FMUL(PULSE_FSUB(o, e), pot);      // :327
  else if (a >= 2) s = PULSE_FMUL(PULSE_FSUB(e, fair), pot);   // :328
  float x = PULSE_FDIV(PULSE_FADD(PULSE_FMUL(pr.w1, m), PULSE_FMUL(pr.w2, s)), (float)pr.K);
  return PULSE_FMUL((float)pr.alpha, tanhf(x));               // :329
}

// get_obs row (:159-179); o must hold 13 + 3*(M-1) floats
template <int MAXP>
PULSE_HD void pk_obs(const PokerTable<MAXP>& t, const PokerParams& pr, float* o) {
  const int AP = pr.AP, W = 13 + 3 * (pr.M - 1);
  const int idx = t.idx;
#pragma unroll
  for (int i = 0; i < 5; ++i) o[i] = (float)t.board[i];
  o[5] = (float)pk_sel(t.hole0, idx);
  o[6] = (float)pk_sel(t.hole1, idx);
  o[7] = (float)t.stage;
  int rel = idx - t.button;  // (idx - button) % AP, python modulo; both are seats in [0, AP)
  if (rel < 0) rel += AP;
  if (rel >= AP) rel %= AP;
  o[8] = (float)rel;
  o[9] = (float)t.pot;
  o[10] = (float)(t.highest - pk_sel(t.crb, idx));
  o[11] = (float)pk_sel(t.stack, idx);
  o[12] = (float)pk_status(t.st, idx);
  int seat = idx;
  for (int k = 1; k < pr.M; ++k) {
    float a = 0.0f, b = 0.0f, c = 0.0f;
    if (k < AP) {
      seat = seat + 1 == AP ? 0 : seat + 1;
      a = (float)pk_sel(t.stack, seat); b = (float)pk_status(t.st, seat); c = (float)pk_sel(t.crb, seat);
    }
    int base = 13 + 3 * (k - 1);
    if (base + 2 < W) { o[base] = a; o[base + 1] = b; o[base + 2] = c; }
  }
}
