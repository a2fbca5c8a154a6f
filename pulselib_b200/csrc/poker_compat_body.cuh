// Per-table bodies of the reference-shaped PokerGPU kernels (see poker_compat.cu).  __host__
// __device__ so tests/hostemu can run the same code on the CPU against the oracle.
#pragma once
#include "poker_core.cuh"

struct CompatDeck {
  const int32_t* row;  // this table's 52 cards
  PULSE_HD int card(int pos) const { return ((unsigned)pos < 52u) ? row[pos] : 0; }
};

template <int MAXP>
PULSE_HD void compat_load(PokerTable<MAXP>& t, const PulsePokerState& S, int g, int P, bool done) {
  uint32_t st = 0xFFFFFFFFu;  // seats beyond P read as SITOUT
#pragma unroll
  for (int s = 0; s < MAXP; ++s) {
    t.stack[s] = t.crb[s] = t.inv[s] = 0; t.hole0[s] = t.hole1[s] = -1;
    if (s < P) {
      size_t o = (size_t)g * P + s;
      t.stack[s] = S.stacks[o]; t.crb[s] = S.current_round_bet[o]; t.inv[s] = S.total_invested[o];
      PulseI2 h = pulse_ld_int2(S.hands + 2 * o);
      t.hole0[s] = h.x; t.hole1[s] = h.y;
      st = pk_set_status(st, s, S.status[o] & 3);
    }
  }
  t.st = st;
  t.pot = S.pots[g]; t.highest = S.highest[g]; t.last_raise = S.last_raise_size[g];
#pragma unroll
  for (int i = 0; i < 5; ++i) t.board[i] = S.board[(size_t)g * 5 + i];
  t.stage = S.stages[g]; t.idx = S.idx[g]; t.agg = S.agg[g]; t.acted = S.acted[g];
  t.button = S.button[g]; t.deck_pos = S.deck_positions[g]; t.eq_stage = 0;
  t.done = done; t.dirty = S.equity_dirty[g] != 0;
}

// writes back only what changed, so untouched tensors cost no store traffic
template <int MAXP>
PULSE_HD void compat_store(const PokerTable<MAXP>& t, const PokerTable<MAXP>& o,
                                             const PulsePokerState& S, int g, int P) {
#pragma unroll
  for (int s = 0; s < MAXP; ++s)
    if (s < P) {
      size_t off = (size_t)g * P + s;
      if (t.stack[s] != o.stack[s]) S.stacks[off] = t.stack[s];
      if (t.crb[s] != o.crb[s]) S.current_round_bet[off] = t.crb[s];
      if (t.inv[s] != o.inv[s]) S.total_invested[off] = t.inv[s];
      int a = pk_status(t.st, s);
      if (a != pk_status(o.st, s)) S.status[off] = a;
    }
  if (t.pot != o.pot) S.pots[g] = t.pot;
  if (t.highest != o.highest) S.highest[g] = t.highest;
  if (t.last_raise != o.last_raise) S.last_raise_size[g] = t.last_raise;
#pragma unroll
  for (int i = 0; i < 5; ++i)
    if (t.board[i] != o.board[i]) S.board[(size_t)g * 5 + i] = t.board[i];
  if (t.stage != o.stage) S.stages[g] = t.stage;
  if (t.idx != o.idx) S.idx[g] = t.idx;
  if (t.agg != o.agg) S.agg[g] = t.agg;
  if (t.acted != o.acted) S.acted[g] = t.acted;
  if (t.deck_pos != o.deck_pos) S.deck_positions[g] = t.deck_pos;
  if (t.dirty != o.dirty) S.equity_dirty[g] = t.dirty;
}

template <int MAXP>
PULSE_HD void compat_step_one(const PulsePokerState& S, const PokerParams& pr, const int64_t* __restrict__ actions,
                              const int32_t* __restrict__ actor_idx, uint32_t phases, float* __restrict__ obs_out,
                              float* __restrict__ reward_out, uint8_t* __restrict__ done_out, int g) {
  const int P = pr.P, AP = pr.AP;
  PokerTable<MAXP> t;
  const bool prev_done = S.is_done[g] != 0;
  compat_load(t, S, g, P, prev_done);
  const PokerTable<MAXP> t0 = t;
  CompatDeck deck{S.decks + (size_t)g * 52};
  const long long a = actions ? actions[g] : 1;
  int actor = t.idx;                                        // :531
  bool legal = pk_status(t.st, actor) == PK_ACTIVE && !prev_done;  // :532-537
  int prev_inv = pk_sel(t.crb, actor);                      // :539
  if (phases & PK_PH_RAW_REWARD) {
    actor = actor_idx ? actor_idx[g] : t.idx;
    prev_inv = S.prev_invested[g];
  } else if ((phases & PK_PH_EXECUTE) && (phases & PK_PH_ADVANCE)) {
    S.prev_invested[g] = prev_inv;                          // written by step (:539), not by execute_actions
  }
  if ((phases & PK_PH_EQUITY) && t.dirty) {                 // :542-543
    for (int s = 0; s < AP; ++s)
      S.equities[(size_t)g * AP + s] =
          pk_equity(t.stage, pk_sel(t.hole0, s), pk_sel(t.hole1, s), t.board, S.hand_ranks, S.n_hand_ranks);
    t.dirty = false;
  }
  if (phases & PK_PH_EXECUTE) pk_execute(t, a);             // :546
  if (phases & PK_PH_ADVANCE) pk_advance(t, pr, deck);      // :547-616
  const bool newly = (phases & PK_PH_NEWLY_IS_DONE) ? t.done : (t.done && !prev_done);  // :619-620
  if ((phases & PK_PH_FOLDWIN) && newly) pk_fold_winner(t, pr);   // :621
  if ((phases & PK_PH_SHOWDOWN) && newly) pk_showdown(t, pr, deck, PokerFaithfulRanker{S.hand_ranks, (long long)S.n_hand_ranks});  // :622
  if ((phases & PK_PH_CLEAR) && t.done) {                   // :625-628
#pragma unroll
    for (int s = 0; s < MAXP; ++s) { t.crb[s] = 0; t.inv[s] = 0; }
    t.highest = 0;
  }
  if (phases & PK_PH_REWARD) {                              // :631-632
    float e = (actor >= 0 && actor < AP) ? S.equities[(size_t)g * AP + actor] : 0.5f;
    float r = pk_reward(t, pr, a, e, prev_inv);
    if (!(phases & PK_PH_RAW_REWARD) && (!legal || prev_done)) r = 0.0f;
    reward_out[g] = r;
  }
  compat_store(t, t0, S, g, P);
  if (done_out && (phases & (PK_PH_ADVANCE))) done_out[g] = t.done;
  if (phases & PK_PH_OBS) {                                 // :633
    float o[13 + 3 * 15];
    pk_obs(t, pr, o);
    const int W = 13 + 3 * (pr.M - 1);
    float* dst = obs_out + (size_t)g * W;
    for (int i = 0; i < W; ++i) dst[i] = o[i];
  }
}

template <int MAXP>
PULSE_HD void compat_reset_one(const PulsePokerState& S, const PokerParams& pr, const int32_t* __restrict__ prefixed,
                               uint64_t seed, uint64_t env_base, uint32_t hand, int rotation, int first_reset,
                               int starting_bbs, int max_bbs, float* __restrict__ obs_out, int g) {
  const int P = pr.P, AP = pr.AP;
  int32_t* deck = S.decks + (size_t)g * 52;
  if (prefixed) {
    for (int i = 0; i < 52; ++i) deck[i] = prefixed[(size_t)g * 52 + i];  // :88-92
  } else {  // chance tape -> forward Fisher-Yates (replaces rand().argsort()+1, :86)
    uint8_t d[52];
    pk_tape_deck(d, seed, env_base + (uint64_t)g, hand);
    for (int i = 0; i < 52; ++i) deck[i] = d[i];
  }
  PokerTable<MAXP> t;
  int32_t old[MAXP];
#pragma unroll
  for (int s = 0; s < MAXP; ++s) {
    old[s] = starting_bbs;
    if (s < P && !(first_reset & 1)) {
      int v = S.stacks[(size_t)g * P + s];
      old[s] = (v == 0 || v > max_bbs) ? starting_bbs : v;  // :105-108
    }
  }
  uint32_t st = 0xFFFFFFFFu;
#pragma unroll
  for (int s = 0; s < MAXP; ++s) {
    int src = s - rotation;  // torch.roll: new[(j + r) % P] = old[j]  (:109-110)
    src %= P; if (src < 0) src += P;
    t.stack[s] = (s < P) ? pk_sel(old, (first_reset & 1) ? s : src) : 0;
    t.crb[s] = 0; t.inv[s] = 0;
    bool dealt = s < AP;
    t.hole0[s] = dealt ? deck[2 * s] : -1;                  // :112-114
    t.hole1[s] = dealt ? deck[2 * s + 1] : -1;
    if (s < P) st = pk_set_status(st, s, dealt ? PK_ACTIVE : PK_SITOUT);  // :118-119
  }
  t.st = st;
  int button = (first_reset & 2) ? 0 : (S.button[g] + 1) % AP;    // :121
  int sb, bb;
  if (AP == 2) { sb = button; bb = (button + 1) % AP; }     // :123-125
  else { sb = (button + 1) % AP; bb = (button + 2) % AP; }  // :127-128
  int bstack = pk_sel(t.stack, bb) - 1;                     // post_blinds :188-199
  pk_put(t.stack, bb, bstack); pk_put(t.crb, bb, 1); pk_put(t.inv, bb, 1);
  if (bstack == 0) t.st = pk_set_status(t.st, bb, PK_ALLIN);
  t.pot = 1; t.highest = 1; t.last_raise = 1;
#pragma unroll
  for (int i = 0; i < 5; ++i) t.board[i] = -1;
  t.stage = 0; t.idx = AP == 2 ? button : (bb + 1) % AP;    // :130-133
  t.agg = bb; t.acted = 0; t.button = button; t.deck_pos = 2 * AP; t.eq_stage = 0; t.done = false; t.dirty = true;
#pragma unroll
  for (int s = 0; s < MAXP; ++s)
    if (s < P) {
      size_t off = (size_t)g * P + s;
      S.stacks[off] = t.stack[s]; S.current_round_bet[off] = t.crb[s]; S.total_invested[off] = t.inv[s];
      S.status[off] = pk_status(t.st, s);
      S.hands[2 * off] = t.hole0[s]; S.hands[2 * off + 1] = t.hole1[s];
    }
  for (int i = 0; i < 5; ++i) S.board[(size_t)g * 5 + i] = -1;
  S.pots[g] = 1; S.stages[g] = 0; S.highest[g] = 1; S.last_raise_size[g] = 1;
  S.idx[g] = t.idx; S.agg[g] = bb; S.acted[g] = 0; S.button[g] = button; S.sb[g] = sb; S.bb[g] = bb;
  S.deck_positions[g] = 2 * AP; S.prev_invested[g] = 0; S.is_done[g] = 0; S.equity_dirty[g] = 1;
  for (int s = 0; s < AP; ++s) S.equities[(size_t)g * AP + s] = 0.5f;  // :144
  if (obs_out) {
    float o[13 + 3 * 15];
    pk_obs(t, pr, o);
    const int W = 13 + 3 * (pr.M - 1);
    for (int i = 0; i < W; ++i) obs_out[(size_t)g * W + i] = o[i];
  }
}
