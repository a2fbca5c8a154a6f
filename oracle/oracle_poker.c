/* TEST INFRASTRUCTURE ONLY -- see oracle_common.h.
 * Scalar, per-table restatement of /root/reference/environments/Poker/PokerGPU.py.  Each function
 * names the reference lines it follows.  The reference applies every tensor op to the whole batch;
 * tables never read each other's state, so running one table through the whole step and then the
 * next gives the same result. */
#include "oracle_poker.h"
#include <limits.h>
#include <math.h>
#include <string.h>
#include <stdlib.h>

enum { ACTIVE = 0, FOLDED = 1, ALLIN = 2, SITOUT = 3 }; /* PokerGPU.py:11 */

static const float RAISE_FRACTIONS[9] = {0.25f, 0.33f, 0.50f, 0.75f, 1.00f, 1.50f, 2.00f, 3.00f, 4.00f}; /* :39 */

/* tape -> deck: forward Fisher-Yates over [1..52]; draw i is lane i%4 of block i/4 of hand `hand` */
void oracle_poker_tape_deck(uint64_t seed, uint64_t env_id, uint32_t hand, int32_t* deck) {
  uint32_t blk[4];
  for (int i = 0; i < 52; ++i) deck[i] = i + 1;
  for (int i = 0; i < 51; ++i) {
    if ((i & 3) == 0) oracle_tape(seed, env_id, ORACLE_STREAM_POKER_DECK, hand * 16u + (uint32_t)(i >> 2), blk);
    int j = i + (int)oracle_mulhi32(blk[i & 3], (uint32_t)(52 - i));
    int32_t t = deck[i]; deck[i] = deck[j]; deck[j] = t;
  }
}

static int live_count(const OraclePoker* s, int g) { /* #ACTIVE + #ALLIN over all P seats */
  int c = 0;
  for (int p = 0; p < s->n_players; ++p) {
    int st = s->status[g * s->n_players + p];
    c += (st == ACTIVE) | (st == ALLIN);
  }
  return c;
}

/* get_obs, PokerGPU.py:159-179 */
void oracle_poker_get_obs(const OraclePoker* s, float* obs) {
  const int P = s->n_players, AP = s->active_players, W = 13 + (s->max_players - 1) * 3;
  for (int g = 0; g < s->n_games; ++g) {
    float* o = obs + (size_t)g * W;
    int idx = s->idx[g];
    for (int i = 0; i < 5; ++i) o[i] = (float)s->board[g * 5 + i];
    o[5] = (float)s->hands[(g * P + idx) * 2];
    o[6] = (float)s->hands[(g * P + idx) * 2 + 1];
    o[7] = (float)s->stages[g];
    int rel = (idx - s->button[g]) % AP; /* python modulo: non-negative */
    if (rel < 0) rel += AP;
    o[8] = (float)rel;
    o[9] = (float)s->pots[g];
    o[10] = (float)(s->highest[g] - s->current_round_bet[g * P + idx]);
    o[11] = (float)s->stacks[g * P + idx];
    o[12] = (float)s->status[g * P + idx];
    for (int i = 13; i < W; ++i) o[i] = 0.0f;
    for (int k = 1; k < AP; ++k) {
      int seat = (idx + k) % AP;
      o[13 + (k - 1) * 3] = (float)s->stacks[g * P + seat];
      o[14 + (k - 1) * 3] = (float)s->status[g * P + seat];
      o[15 + (k - 1) * 3] = (float)s->current_round_bet[g * P + seat];
    }
  }
}

/* reset, PokerGPU.py:73-157.  `decks` must already hold the [N,52] decks (prefixed or from the
 * tape).  first_reset: stacks are filled and the button starts at 0; otherwise busted / oversized
 * stacks are refilled, rolled by `rotation`, and the button advances modulo the new table size. */
void oracle_poker_reset(OraclePoker* s, int active_players, int first_reset, int rotation) {
  const int N = s->n_games, P = s->n_players;
  s->active_players = active_players;
  const int AP = active_players;
  for (int g = 0; g < N; ++g) {
    s->last_raise_size[g] = 1;                                  /* :81 */
    s->deck_positions[g] = 0;                                   /* :93 */
    for (int i = 0; i < 5; ++i) s->board[g * 5 + i] = -1;       /* :95 */
    s->pots[g] = 0; s->stages[g] = 0;                           /* :96-97 */
    int32_t* st = s->stacks + (size_t)g * P;
    if (first_reset) {
      for (int p = 0; p < P; ++p) st[p] = s->starting_bbs;      /* :102 */
    } else {
      for (int p = 0; p < P; ++p)
        if (st[p] == 0 || st[p] > s->max_bbs) st[p] = s->starting_bbs; /* :105-108 */
      if (rotation != 0) {                                      /* :109-110 torch.roll */
        int32_t tmp[64];
        for (int p = 0; p < P; ++p) {
          int q = (p + rotation) % P;
          if (q < 0) q += P;
          tmp[q] = st[p];
        }
        memcpy(st, tmp, sizeof(int32_t) * P);
      }
    }
    for (int p = 0; p < P; ++p) {                               /* :112-119 */
      int dealt = p < AP;
      s->hands[(g * P + p) * 2] = dealt ? s->decks[g * 52 + 2 * p] : -1;
      s->hands[(g * P + p) * 2 + 1] = dealt ? s->decks[g * 52 + 2 * p + 1] : -1;
      s->current_round_bet[g * P + p] = 0;
      s->total_invested[g * P + p] = 0;
      s->status[g * P + p] = dealt ? ACTIVE : SITOUT;
    }
    s->deck_positions[g] = 2 * AP;                              /* :205 */
    s->button[g] = first_reset ? 0 : (s->button[g] + 1) % AP;   /* :121 */
    if (AP == 2) { s->sb[g] = s->button[g]; s->bb[g] = (s->button[g] + 1) % AP; }       /* :123-125 */
    else { s->sb[g] = (s->button[g] + 1) % AP; s->bb[g] = (s->button[g] + 2) % AP; }    /* :127-128 */
    /* post_blinds :188-199 -- big blind only, one chip */
    int b = s->bb[g];
    st[b] -= 1;
    s->current_round_bet[g * P + b] += 1;
    s->total_invested[g * P + b] += 1;
    s->pots[g] += 1;
    s->status[g * P + b] = st[b] == 0 ? ALLIN : ACTIVE;
    s->idx[g] = AP == 2 ? s->button[g] : (s->bb[g] + 1) % AP;   /* :130-133 */
    s->highest[g] = 1; s->agg[g] = s->bb[g]; s->acted[g] = 0;   /* :134-136 */
    s->is_done[g] = 0;                                          /* :137 */
    s->prev_invested[g] = 0;                                    /* :151 */
    s->equity_dirty[g] = 1;                                     /* :145 */
    for (int p = 0; p < AP; ++p) s->equities[g * AP + p] = 0.5f; /* :144 */
  }
}

static int hr_at(const OraclePoker* s, int64_t i) {
  /* torch wraps negative indices; out-of-range would raise in the reference -- the oracle returns
   * 0 there so a stub table (tests/poker/test_poker_gpu_prefixed_decks.py:14-34) stays walkable */
  if (i < 0) i += s->n_hand_ranks;
  if (i < 0 || i >= s->n_hand_ranks) return 0;
  return s->hand_ranks[i];
}

/* calculate_equities for one table, PokerGPU.py:455-525 */
static void calc_equities_table(OraclePoker* s, int g) {
  const int P = s->n_players, AP = s->active_players;
  if (!s->equity_dirty[g]) return;
  int stage = s->stages[g];
  for (int p = 0; p < AP; ++p) {
    float e = 0.5f;                                             /* :461 */
    int ncards = stage == 3 ? 7 : stage == 2 ? 6 : stage == 1 ? 5 : 0;
    if (ncards) {
      int off = 53;
      for (int c = 0; c < ncards; ++c) {
        int card = c < 2 ? s->hands[(g * P + p) * 2 + c] : s->board[g * 5 + (c - 2)];
        off = hr_at(s, (int64_t)off + card);                    /* :477-478, :497-498, :518-519 */
      }
      float v, lo, den;
      if (stage == 3) { v = (float)off; lo = 4109.0f; den = 32765.0f; }                    /* :480-481 */
      else if (stage == 2) { v = (float)hr_at(s, off); lo = 4109.0f; den = 32765.0f; }     /* :500-502 */
      else { v = (float)hr_at(s, hr_at(s, off)); lo = 74359.0f; den = 749420.0f; }         /* :521-523 */
      e = (v - lo) / den;
      e = e < 0.0f ? 0.0f : (e > 1.0f ? 1.0f : e);
    }
    s->equities[g * AP + p] = e;
  }
  s->equity_dirty[g] = 0;                                       /* :525 */
}

void oracle_poker_calculate_equities(OraclePoker* s) {
  for (int g = 0; g < s->n_games; ++g) calc_equities_table(s, g);
}

/* execute_actions for one table, PokerGPU.py:230-303 */
static void execute_action_table(OraclePoker* s, int g, int64_t a) {
  const int P = s->n_players;
  int idx = s->idx[g];
  int32_t* stack = &s->stacks[g * P + idx];
  int32_t* crb = &s->current_round_bet[g * P + idx];
  int32_t* inv = &s->total_invested[g * P + idx];
  int32_t* stat = &s->status[g * P + idx];
  int call_cost = s->highest[g] - *crb;                         /* :232 */
  int active = *stat != FOLDED && *stat != ALLIN && *stat != SITOUT && !s->is_done[g]; /* :233 */
  if (!active) return;
  if (a == 0) {                                                 /* :236-238 */
    *stat = FOLDED;
    s->acted[g] += 1;
  } else if (a == 1) {                                          /* :241-252 */
    int amt = call_cost < *stack ? call_cost : *stack;
    *stack -= amt; *crb += amt; *inv += amt; s->pots[g] += amt;
    if (*stack == 0) *stat = ALLIN;
    s->acted[g] += 1;
  } else if (a >= 2) {                                          /* :255-303 */
    int raise_amt = 0;
    if (a == 2) raise_amt = s->last_raise_size[g];              /* :258-259 */
    else if (a == 12) raise_amt = *stack;                       /* :262-264 */
    else if (a >= 3 && a <= 11) {
      volatile float prod = (float)s->pots[g] * RAISE_FRACTIONS[a - 3]; /* :270 fp32 multiply */
      raise_amt = (int32_t)prod;                                /* .to(int32) truncates */
    }
    int total = call_cost + raise_amt;                          /* :272 */
    int bet = total < *stack ? total : *stack;                  /* :273 */
    int is_raise = !(bet <= call_cost);                         /* :274-275 */
    *stack -= bet; *crb += bet; *inv += bet; s->pots[g] += bet; /* :277-280 */
    if (*stack == 0) *stat = ALLIN;                             /* :282-286 */
    if (is_raise) {                                             /* :288-301 */
      int new_bet = *crb, old_high = s->highest[g];
      int size = new_bet - old_high;
      s->highest[g] = new_bet;
      if (size >= s->last_raise_size[g]) {
        s->agg[g] = idx; s->acted[g] = 0; s->last_raise_size[g] = size;
      }
    }
    s->acted[g] += 1;                                           /* :303 */
  }
  /* negative action ids match no mask in the reference: nothing happens */
}

void oracle_poker_execute_actions(OraclePoker* s, const int64_t* actions) {
  for (int g = 0; g < s->n_games; ++g) execute_action_table(s, g, actions[g]);
}

/* deal_cards for one table, PokerGPU.py:208-214 */
static int32_t deal_one(OraclePoker* s, int g) {
  int pos = s->deck_positions[g];
  int32_t c = (pos >= 0 && pos < 52) ? s->decks[g * 52 + pos] : 0;
  s->deck_positions[g] = pos + 1;
  return c;
}

/* _award_showdown_side_pots for one table, PokerGPU.py:340-378 */
static void award_side_pots_table(OraclePoker* s, int g, const int32_t* ranks, const int* eligible) {
  const int P = s->n_players, AP = s->active_players;
  int32_t inv[64], sorted[64], payout[64];
  for (int p = 0; p < AP; ++p) { inv[p] = s->total_invested[g * P + p]; sorted[p] = inv[p]; payout[p] = 0; }
  for (int i = 1; i < AP; ++i) {                                /* :347 ascending sort */
    int32_t v = sorted[i]; int j = i - 1;
    while (j >= 0 && sorted[j] > v) { sorted[j + 1] = sorted[j]; --j; }
    sorted[j + 1] = v;
  }
  for (int k = 0; k < AP; ++k) {
    int32_t size = sorted[k] - (k ? sorted[k - 1] : 0);         /* :348-352 */
    int contributors = 0, winners = 0, first_winner = 0;
    int32_t best = INT_MIN;
    for (int p = 0; p < AP; ++p) {
      int contrib = inv[p] >= sorted[k];                        /* :353 */
      contributors += contrib;
      if (contrib && eligible[p] && ranks[p] > best) best = ranks[p]; /* :354-357 */
    }
    for (int p = AP - 1; p >= 0; --p) {
      int w = (inv[p] >= sorted[k]) && eligible[p] && ranks[p] == best; /* :358 */
      if (w) { winners++; first_winner = p; }
    }
    int32_t pot = size * contributors;                          /* :359-360 */
    if (!(size > 0 && winners > 0)) continue;                   /* :362 */
    int32_t share = pot / winners, rem = pot % winners;         /* :364-373 (pot >= 0: floor == trunc) */
    for (int p = 0; p < AP; ++p)
      if ((inv[p] >= sorted[k]) && eligible[p] && ranks[p] == best) payout[p] += share; /* :374 */
    payout[first_winner] += rem;                                /* :375-376 lowest-index winner */
  }
  for (int p = 0; p < AP; ++p) s->stacks[g * P + p] += payout[p]; /* :377-378 */
}

/* resolve_fold_winners for one table; `newly` is the temporary is_done of :620.  :331-338 */
static void resolve_fold_winner_table(OraclePoker* s, int g, int newly) {
  const int P = s->n_players;
  if (!newly || live_count(s, g) != 1) return;
  for (int p = 0; p < P; ++p) {
    int st = s->status[g * P + p];
    if (st == ACTIVE || st == ALLIN) { s->stacks[g * P + p] += s->pots[g]; break; }
  }
  s->pots[g] = 0;
}

/* resolve_terminated_games for one table, PokerGPU.py:380-453 */
static void resolve_terminated_table(OraclePoker* s, int g, int newly) {
  const int P = s->n_players, AP = s->active_players;
  if (!(s->stages[g] < 5 && newly)) return;                     /* :386 */
  if (!(live_count(s, g) > 1)) return;                          /* :390 */
  int stage = s->stages[g];
  if (stage == 0) {                                             /* :397-405 */
    s->deck_positions[g] += 1;
    for (int i = 0; i < 3; ++i) s->board[g * 5 + i] = deal_one(s, g);
    s->deck_positions[g] += 1; s->board[g * 5 + 3] = deal_one(s, g);
    s->deck_positions[g] += 1; s->board[g * 5 + 4] = deal_one(s, g);
  } else if (stage == 1) {                                      /* :407-414 */
    s->deck_positions[g] += 1; s->board[g * 5 + 3] = deal_one(s, g);
    s->deck_positions[g] += 1; s->board[g * 5 + 4] = deal_one(s, g);
  } else if (stage == 2) {                                      /* :416-419 */
    s->deck_positions[g] += 1; s->board[g * 5 + 4] = deal_one(s, g);
  }
  int32_t ranks[64]; int eligible[64];
  for (int p = 0; p < AP; ++p) {                                /* :427-450 */
    int off = 53;
    for (int c = 0; c < 7; ++c) {
      int card = c < 2 ? s->hands[(g * P + p) * 2 + c] : s->board[g * 5 + (c - 2)];
      off = hr_at(s, (int64_t)off + card);
    }
    int st = s->status[g * P + p];
    eligible[p] = (st == ACTIVE) || (st == ALLIN);
    ranks[p] = eligible[p] ? off : INT_MIN;
  }
  award_side_pots_table(s, g, ranks, eligible);                 /* :451 */
  s->pots[g] = 0; s->stages[g] = 5;                             /* :452-453 */
}

void oracle_poker_resolve_fold_winners(OraclePoker* s) { /* as called directly: is_done is the mask */
  for (int g = 0; g < s->n_games; ++g) resolve_fold_winner_table(s, g, s->is_done[g]);
}
void oracle_poker_resolve_terminated_games(OraclePoker* s) {
  for (int g = 0; g < s->n_games; ++g) resolve_terminated_table(s, g, s->is_done[g]);
}

/* poker_reward_gpu for one table, PokerGPU.py:305-329 (fp32 throughout, one rounding per op) */
static float reward_table(const OraclePoker* s, int g, int64_t a, int actor) {
  const int AP = s->active_players;
  volatile float cnt = (float)live_count(s, g);                 /* :309 */
  volatile float fair = 1.0f / (cnt < 1.0f ? 1.0f : cnt);       /* :310 */
  int cc = s->highest[g] - s->prev_invested[g];                 /* :311 */
  if (cc < 0) cc = 0;
  float e = s->equities[g * AP + actor];                        /* :314 */
  volatile float pot = (float)s->pots[g];
  volatile float m = e * pot;                                   /* :315 */
  volatile float den = (float)(s->pots[g] + cc) + 1e-6f;        /* :316 */
  volatile float o = (float)cc / den;
  volatile float sv = 0.0f, d;                                  /* :306 */
  if (a == 1) { d = e - o; sv = d * pot; }                      /* :326 */
  if (a == 0) { d = o - e; sv = d * pot; }                      /* :327 */
  if (a >= 2) { d = e - fair; sv = d * pot; }                   /* :328 */
  volatile float t1 = s->w1 * m, t2 = s->w2 * sv;
  volatile float x = (t1 + t2) / (float)s->K;                   /* :329 */
  volatile float th = tanhf(x);
  return (float)s->alpha * th;
}

/* poker_reward_gpu called on its own (PokerGPU.py:305-329): no legality masking, actor given by the caller */
void oracle_poker_reward(const OraclePoker* s, const int64_t* actions, const int32_t* actor_idx, float* out) {
  for (int g = 0; g < s->n_games; ++g) out[g] = reward_table(s, g, actions[g], actor_idx[g]);
}

/* step, PokerGPU.py:527-633.  use_cached_equities != 0 skips step 1 (a caller that replaced
 * calculate_equities has already run its own). */
void oracle_poker_step(OraclePoker* s, const int64_t* actions, float* obs, float* rewards,
                       int use_cached_equities) {
  const int N = s->n_games, P = s->n_players, AP = s->active_players;
#pragma omp parallel for if (N > 4096) schedule(static)
  for (int g = 0; g < N; ++g) {
    int prev_done = s->is_done[g];                              /* :530 */
    int actor = s->idx[g];                                      /* :531 */
    int ast = s->status[g * P + actor];
    int legal = ast != FOLDED && ast != ALLIN && ast != SITOUT && !prev_done; /* :532-537 */
    s->prev_invested[g] = s->current_round_bet[g * P + actor];  /* :539 */
    if (!use_cached_equities) calc_equities_table(s, g);        /* :542-543 */
    execute_action_table(s, g, actions[g]);                     /* :546 */
    int truly_active = 0;                                       /* :547 */
    for (int p = 0; p < P; ++p) truly_active += s->status[g * P + p] == ACTIVE;
    int all_acted = s->acted[g] >= truly_active;                /* :549 */
    int round_over = s->is_done[g] || truly_active == 0;        /* :552-553 */
    int has_next = 0, next_seat = 0;                            /* :554-561 */
    for (int k = 1; k <= AP; ++k) {
      int seat = (s->idx[g] + k) % AP;
      if (s->status[g * P + seat] == ACTIVE) { has_next = 1; next_seat = seat; break; }
    }
    if (!has_next) next_seat = (s->idx[g] + 1) % AP;            /* argmax of all-False = 0 */
    if (!round_over) {                                          /* :562-569 */
      int closes_cur = all_acted && s->idx[g] == s->agg[g];
      int closes_next = all_acted && has_next && next_seat == s->agg[g];
      round_over = !has_next || closes_cur || closes_next;
    }
    if (!round_over && has_next) s->idx[g] = next_seat;         /* :571-573 */
    int lc = live_count(s, g);                                  /* :576 */
    int early = lc <= 1 && round_over;                          /* :577 */
    if (early) s->is_done[g] = 1;                               /* :578 */
    int trans = round_over && !early && !s->is_done[g];         /* :580 */
    if (trans) {
      s->last_raise_size[g] = 1;                                /* :581 */
      s->stages[g] += 1;                                        /* :584 */
      s->highest[g] = 0;                                        /* :585 */
      s->agg[g] = (s->button[g] + 1) % AP;                      /* :586 */
      s->acted[g] = 0;                                          /* :587 */
      for (int p = 0; p < P; ++p) s->current_round_bet[g * P + p] = 0; /* :588 */
      for (int k = 1; k <= AP; ++k) {                           /* :589 -> :215-228 */
        int seat = (s->button[g] + k) % AP;
        if (s->status[g * P + seat] == ACTIVE) { s->idx[g] = seat; break; }
      }
      int stage = s->stages[g];
      if (stage > 3) { s->is_done[g] = 1; s->stages[g] = 4; }   /* :595-598 */
      else if (stage == 1) {                                    /* :601-604 */
        s->deck_positions[g] += 1;
        for (int i = 0; i < 3; ++i) s->board[g * 5 + i] = deal_one(s, g);
        s->equity_dirty[g] = 1;
      } else if (stage == 2) {                                  /* :607-610 */
        s->deck_positions[g] += 1; s->board[g * 5 + 3] = deal_one(s, g); s->equity_dirty[g] = 1;
      } else if (stage == 3) {                                  /* :613-616 */
        s->deck_positions[g] += 1; s->board[g * 5 + 4] = deal_one(s, g); s->equity_dirty[g] = 1;
      }
    }
    int all_done = s->is_done[g];                               /* :619 */
    int newly = all_done && !prev_done;                         /* :620 */
    resolve_fold_winner_table(s, g, newly);                     /* :621 */
    resolve_terminated_table(s, g, newly);                      /* :622 */
    s->is_done[g] = (uint8_t)all_done;                          /* :623 */
    if (all_done) {                                             /* :625-628 */
      for (int p = 0; p < P; ++p) { s->current_round_bet[g * P + p] = 0; s->total_invested[g * P + p] = 0; }
      s->highest[g] = 0;
    }
    float r = reward_table(s, g, actions[g], actor);            /* :631 */
    if (!legal || prev_done) r = 0.0f;                          /* :632 */
    rewards[g] = r;
  }
  if (obs) oracle_poker_get_obs(s, obs);                        /* :633 */
}

/* fills s->decks from the chance tape for every table (OpenMP) */
void oracle_poker_fill_tape_decks(OraclePoker* s, uint64_t seed, uint64_t env_base, uint32_t hand) {
#pragma omp parallel for schedule(static)
  for (int g = 0; g < s->n_games; ++g) oracle_poker_tape_deck(seed, env_base + (uint64_t)g, hand, s->decks + (size_t)g * 52);
}

/* actions from the chance tape: uniform over 0..n_actions-1, stream ACTION, counter = step */
void oracle_tape_actions(uint64_t seed, uint64_t env_base, uint32_t step, int n_actions, int64_t* out, int n) {
#pragma omp parallel for schedule(static)
  for (int g = 0; g < n; ++g) {
    uint32_t blk[4];
    oracle_tape(seed, env_base + (uint64_t)g, ORACLE_STREAM_ACTION, step, blk);
    out[g] = (int64_t)oracle_mulhi32(blk[0], (uint32_t)n_actions);
  }
}

/* ---- scripted opponents + build_actions -------------------------------------------
 * Restates /root/reference/environments/Poker/Player.py:79-176 (the four scripted GPU agents) and the
 * router /root/reference/environments/Poker/utils.py:108-123, one table at a time, in the reference's
 * own order of mask assignments (default action, fold mask, call mask, raise mask -- later masks
 * overwrite earlier ones exactly as the tensor assignments do).
 * kinds: 0 external (action untouched), 1 random, 2 heuristic_hands, 3 tight_aggressive,
 * 4 loose_passive, 5 small_ball.  chance2 [N,2]: word 0 -> torch.randint(lo, hi) = lo + mulhi32(w0, hi-lo),
 * word 1 -> torch.rand = (w1 >> 8) * 2^-24. */
static int py_mod13(int v) { int r = v % 13; return r < 0 ? r + 13 : r; }
static int tape_randint(uint32_t w, int lo, int hi) { return lo + (int)(((uint64_t)w * (uint64_t)(hi - lo)) >> 32); }

void oracle_poker_policy(const float* obs, int W, const int32_t* seat_idx, const uint8_t* seat_kinds, int n_seats,
                         const uint32_t* chance2, int64_t* actions, int n) {
  static const int raise_distribution[9] = {2, 3, 4, 5, 6, 7, 8, 9, 10};  /* torch.arange(2, 11) */
  for (int g = 0; g < n; ++g) {
    int seat = seat_idx[g];
    int kind = (seat >= 0 && seat < n_seats) ? seat_kinds[seat] : 0;
    if (kind == 0) continue;                                   /* utils.py:117-118: the Q agent's own call */
    uint32_t w0 = chance2[2 * g], w1 = chance2[2 * g + 1];
    if (kind == 1) { actions[g] = tape_randint(w0, 0, 13); continue; }   /* utils.py:119-120 */
    const float* st = obs + (size_t)g * W;
    int rank1 = py_mod13((int)st[5]), rank2 = py_mod13((int)st[6]);    /* hands = states[:, 5:7]; ranks = hands % 13 */
    int diff = abs(rank1 - rank2);
    int64_t act;
    if (kind == 2) {                                           /* HeuristicHandsPlayerGPU.action :86-101 */
      act = 0;
      int fold_mask = rank1 < 8 && rank2 < 8;
      if (fold_mask) act = 0;
      int pair_mask = rank1 == rank2;
      int high_card_mask = rank1 >= 10 || rank2 >= 10;
      int raise_mask = (pair_mask || high_card_mask) && !fold_mask;
      if (raise_mask) act = raise_distribution[tape_randint(w0, 0, 9)];
    } else if (kind == 3) {                                    /* TightAggressivePlayerGPU.action :111-124 */
      act = 1;
      int fold_mask = rank1 < 7 && rank2 < 7 && diff > 5;
      if (fold_mask) act = 0;
      int pair_mask = rank1 == rank2;
      int high_card_mask = (rank1 >= 10 && rank2 > 5) || (rank2 >= 10 && rank1 > 5);
      int raise_mask = (pair_mask || high_card_mask) && !fold_mask;
      if (raise_mask) act = raise_distribution[tape_randint(w0, 5, 9)];
    } else if (kind == 4) {                                    /* LoosePassivePlayerGPU.action :134-148 */
      act = 0;
      float probs = (float)(w1 >> 8) * (1.0f / 16777216.0f);
      int fold_mask = rank1 <= 4 && rank2 <= 4 && diff > 9;
      if (fold_mask) act = 0;
      int pair_mask = rank1 == rank2 && rank1 > 8;
      int high_card_mask = (rank1 >= 11 && rank2 > 9) || (rank2 >= 11 && rank1 > 9);
      int call_mask = (pair_mask || high_card_mask) && !fold_mask;
      int raise_mask = probs > 0.9f && call_mask;
      if (call_mask) act = 1;
      if (raise_mask) act = raise_distribution[tape_randint(w0, 0, 4)];
    } else {                                                   /* SmallBallPlayerGPU.action :158-174 */
      act = 0;
      float pot_size = st[9];
      int fold_mask = (rank1 < 6 && rank2 < 6 && pot_size > 30.0f) || (rank1 < 9 && rank2 < 9 && pot_size > 80.0f);
      if (fold_mask) act = 0;
      int pair_mask = rank1 == rank2;
      int high_card_mask = (rank1 >= 10 && rank2 > 5) || (rank2 >= 10 && rank1 > 5);
      int raise_mask = (pair_mask || high_card_mask) && !fold_mask;
      if (raise_mask) act = raise_distribution[tape_randint(w0, 0, 3)];
    }
    actions[g] = act;
  }
}

/* the tape's chance pair for the policy (stream 2, counter = step): words x, y of the Philox block */
void oracle_policy_tape(uint64_t seed, uint64_t env_base, uint32_t counter, uint32_t* chance2, int n) {
  for (int g = 0; g < n; ++g) {
    uint32_t b[4];
    oracle_tape(seed, env_base + (uint64_t)g, 2u, counter, b);
    chance2[2 * g] = b[0]; chance2[2 * g + 1] = b[1];
  }
}
