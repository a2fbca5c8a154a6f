/* TEST INFRASTRUCTURE ONLY -- see oracle_common.h.
 * CPU restatement of /root/reference/environments/blackjack/blackjack.py (BlackJack.reset :23-48,
 * deal_starting_cards :51-101, execute_actions :113-162, calculate_rewards :164-177, step :179-186,
 * get_obs :103-108) on reference-shaped arrays: one array per tensor attribute of the reference.
 * Parity pinned by tests/golden/blackjack.npz, produced by running the reference file in place
 * (tests/golden/make_golden.py). */
#include "oracle_common.h"

typedef struct {
  int32_t n;                  /* batch_size */
  int32_t* decks;             /* [N,52] permutation of 0..51                 :24-29 */
  int32_t* deck_positions;    /* [N]                                         :31 */
  uint8_t* terminated;        /* [N]                                         :32 */
  int32_t* players_cards;     /* [N,20] counted values (ace as 11 or 1)      :33 */
  int32_t* players_card_idx;  /* [N]                                         :34 */
  int32_t* player_card_sums;  /* [N]                                         :35 */
  uint8_t* has_ace;           /* [N] usable-ace flag                         :36 */
  int32_t* dealer_cards;      /* [N,20]                                      :38 */
  int32_t* dealer_card_idx;   /* [N]                                         :39 */
  int32_t* dealer_upcard;     /* [N]                                         :40 */
  int32_t* dealer_card_sums;  /* [N]                                         :41 */
  uint8_t* dealer_has_ace;    /* [N]                                         :42 */
  int32_t* rewards;           /* [N]                                         :43 */
  int32_t* obs;               /* [N,3]                                       :44 */
} OracleBlackjack;

static int card_value(int card) { /* clamp(card % 13 + 1, max=10)  :53 */
  int v = card % 13 + 1;
  return v > 10 ? 10 : v;
}

/* Fisher-Yates deck of 0..51 from the chance tape, the same construction as the poker deck */
void oracle_blackjack_tape_decks(uint64_t seed, uint64_t env_base, uint32_t episode, int n, int32_t* decks) {
  for (int g = 0; g < n; ++g) {
    int32_t* d = decks + (size_t)g * 52;
    for (int i = 0; i < 52; ++i) d[i] = i;
    uint32_t blk[4] = {0, 0, 0, 0};
    for (int i = 0; i < 52; ++i) {
      if ((i & 3) == 0) oracle_tape(seed, env_base + (uint64_t)g, ORACLE_STREAM_BJ_DECK, episode * 16u + (uint32_t)(i >> 2), blk);
      int j = i + (int)oracle_mulhi32(blk[i & 3], (uint32_t)(52 - i));
      int32_t t = d[i]; d[i] = d[j]; d[j] = t;
    }
  }
}

static void write_obs(OracleBlackjack* s, int g) { /* :103-108 */
  s->obs[3 * g] = s->player_card_sums[g];
  s->obs[3 * g + 1] = s->has_ace[g];
  s->obs[3 * g + 2] = s->dealer_upcard[g];
}

/* reset + deal_starting_cards; s->decks already holds the decks */
void oracle_blackjack_reset(OracleBlackjack* s) {
  for (int g = 0; g < s->n; ++g) {
    const int32_t* d = s->decks + (size_t)g * 52;
    int32_t* pc = s->players_cards + (size_t)g * 20;
    int32_t* dc = s->dealer_cards + (size_t)g * 20;
    for (int i = 0; i < 20; ++i) pc[i] = dc[i] = 0;
    s->terminated[g] = 0; s->rewards[g] = 0;
    /* opening deal: every ace counts 11 (:53-55, :62-64, :72-74, :81-83); order player, dealer, player, dealer */
    int p1 = card_value(d[0]), d1 = card_value(d[1]), p2 = card_value(d[2]), d2 = card_value(d[3]);
    int pa1 = p1 == 1, da1 = d1 == 1, pa2 = p2 == 1, da2 = d2 == 1;
    if (pa1) p1 = 11;
    if (da1) d1 = 11;
    if (pa2) p2 = 11;
    if (da2) d2 = 11;
    pc[0] = p1; pc[1] = p2; dc[0] = d1; dc[1] = d2;
    s->players_card_idx[g] = 2; s->dealer_card_idx[g] = 2; s->deck_positions[g] = 4;
    s->dealer_upcard[g] = d1;                                    /* :67 */
    int ha = pa1 | pa2;                                          /* :58, :78 */
    int dha = da1 | (!da1 && da2);                               /* :69, :82-85 */
    int ps = p1 + p2, ds = d1 + d2;                              /* :89-90 */
    if (ps > 21 && ha) { ps -= 10; ha = 0; }                     /* :93-95 */
    if (ds > 21 && dha) { ds -= 10; dha = 0; }                   /* :99-101 */
    s->player_card_sums[g] = ps; s->dealer_card_sums[g] = ds;
    s->has_ace[g] = (uint8_t)ha; s->dealer_has_ace[g] = (uint8_t)dha;
    write_obs(s, g);
  }
}

/* step (:179-186): execute_actions, rewards.zero_(), calculate_rewards, get_obs */
void oracle_blackjack_step(OracleBlackjack* s, const int64_t* actions) {
#pragma omp parallel for schedule(static) if (s->n > 16384)
  for (int g = 0; g < s->n; ++g) {
    const int32_t* d = s->decks + (size_t)g * 52;
    const int64_t a = actions[g];
    const int hit = a == 0 && !s->terminated[g];                 /* :117 */
    const int stand = a == 1 && !s->terminated[g];               /* :138 */
    if (hit) {
      int r = card_value(d[s->deck_positions[g]]);               /* :118-119 */
      const int first_ace = r == 1 && !s->has_ace[g];            /* :120-122 */
      if (first_ace) r = 11;
      s->players_cards[(size_t)g * 20 + s->players_card_idx[g]] = r;
      s->has_ace[g] |= (uint8_t)first_ace;                       /* :127 */
      s->player_card_sums[g] += r;
      s->deck_positions[g] += 1; s->players_card_idx[g] += 1;
      if (s->player_card_sums[g] > 21 && s->has_ace[g]) {        /* :133-135 */
        s->player_card_sums[g] -= 10; s->has_ace[g] = 0;
      }
    }
    if (stand) {
      int active = s->dealer_card_sums[g] < 17;                  /* :139 */
      while (active) {                                           /* :141-160 */
        int r = card_value(d[s->deck_positions[g]]);
        const int first_ace = r == 1 && !s->dealer_has_ace[g];
        if (first_ace) r = 11;
        s->dealer_cards[(size_t)g * 20 + s->dealer_card_idx[g]] = r;
        s->dealer_card_idx[g] += 1;
        s->dealer_has_ace[g] |= (uint8_t)first_ace;
        s->dealer_card_sums[g] += r;
        if (s->dealer_card_sums[g] > 21 && s->dealer_has_ace[g]) { s->dealer_card_sums[g] -= 10; s->dealer_has_ace[g] = 0; }
        s->deck_positions[g] += 1;
        active = s->dealer_card_sums[g] < 17 && s->dealer_card_sums[g] <= 21;
      }
    }
    s->rewards[g] = 0;                                           /* :183 */
    if (hit && s->player_card_sums[g] > 21) { s->rewards[g] = -1; s->terminated[g] = 1; }  /* :166-168 */
    if (stand) {                                                 /* :171-177 push counts as a win */
      const int ps = s->player_card_sums[g], ds = s->dealer_card_sums[g];
      s->rewards[g] = (ds > 21 || ps >= ds) ? 1 : -1;
      s->terminated[g] = 1;
    }
    write_obs(s, g);
  }
}
