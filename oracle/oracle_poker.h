/* TEST INFRASTRUCTURE ONLY -- see oracle_common.h.
 * Reference-shaped PokerGPU state (environments/Poker/PokerGPU.py:73-157): one int32 array per
 * torch tensor of the reference, same shapes, row-major. */
#ifndef PULSE_ORACLE_POKER_H
#define PULSE_ORACLE_POKER_H
#include "oracle_common.h"

typedef struct {
  int32_t n_games, n_players, active_players, max_players;
  int32_t starting_bbs, max_bbs;
  int32_t K, alpha;
  float w1, w2;
  int32_t* decks;             /* [N,52] values 1..52 */
  int32_t* deck_positions;    /* [N] */
  int32_t* board;             /* [N,5], -1 = not dealt */
  int32_t* pots;              /* [N] */
  int32_t* stages;            /* [N] */
  int32_t* stacks;            /* [N,P] */
  int32_t* hands;             /* [N,P,2] */
  int32_t* current_round_bet; /* [N,P] */
  int32_t* total_invested;    /* [N,P] */
  int32_t* status;            /* [N,P] 0 ACTIVE 1 FOLDED 2 ALLIN 3 SITOUT */
  int32_t* button;            /* [N] */
  int32_t* sb;                /* [N] */
  int32_t* bb;                /* [N] */
  int32_t* idx;               /* [N] */
  int32_t* highest;           /* [N] */
  int32_t* agg;               /* [N] */
  int32_t* acted;             /* [N] */
  int32_t* last_raise_size;   /* [N] */
  int32_t* prev_invested;     /* [N] */
  uint8_t* is_done;           /* [N] */
  uint8_t* equity_dirty;      /* [N] */
  float* equities;            /* [N,AP] */
  const int32_t* hand_ranks;  /* HandRanks table */
  int64_t n_hand_ranks;
} OraclePoker;

#endif
