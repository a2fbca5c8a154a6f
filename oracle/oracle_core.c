/* TEST INFRASTRUCTURE ONLY -- see oracle_common.h.
 * Chance tape (Philox), an independent brute-force poker hand evaluator used to validate the
 * regenerated HandRanks table, and the Particle2D restatement. */
#include "oracle_common.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

/* ---- Philox-4x32-10 (Salmon et al., SC'11; Random123 constants) ------------------------------ */
void oracle_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
  uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3], k0 = key[0], k1 = key[1];
  for (int r = 0; r < 10; ++r) {
    uint64_t p0 = (uint64_t)0xD2511F53u * c0, p1 = (uint64_t)0xCD9E8D57u * c2;
    uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
    uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
    c1 = (uint32_t)p1; c3 = (uint32_t)p0; c0 = n0; c2 = n2;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

void oracle_tape(uint64_t seed, uint64_t env_id, uint32_t stream, uint32_t counter, uint32_t out[4]) {
  uint32_t ctr[4] = {(uint32_t)env_id, (uint32_t)(env_id >> 32), counter, stream};
  uint32_t key[2] = {(uint32_t)seed, (uint32_t)(seed >> 32)};
  oracle_philox4x32_10(ctr, key, out);
}

/* ---- brute-force evaluator -------------------------------------------------------------------
 * Strength key of five cards (cards 1..52, rank = (c-1)>>2, suit = (c-1)&3): category in the top
 * nibble then tie-break ranks, most significant first.  The 7462 distinct keys, sorted, give
 * rank-in-category; value = (category << 12) + rank, the HandRanks convention. */
static uint32_t key5(const int c[5]) {
  int r[5], cnt[13] = {0};
  int flush = 1;
  for (int i = 0; i < 5; ++i) {
    r[i] = (c[i] - 1) >> 2;
    cnt[r[i]]++;
    if (((c[i] - 1) & 3) != ((c[0] - 1) & 3)) flush = 0;
  }
  /* order ranks by (count desc, rank desc) */
  int ord[5], no = 0;
  for (int want = 4; want >= 1; --want)
    for (int rk = 12; rk >= 0; --rk)
      if (cnt[rk] == want) ord[no++] = rk;
  int distinct = no;
  int straight = 0, top = 0;
  if (distinct == 5) {
    if (ord[0] - ord[4] == 4) { straight = 1; top = ord[0]; }
    else if (ord[0] == 12 && ord[1] == 3 && ord[4] == 0) { straight = 1; top = 3; }
  }
  int cat;
  if (straight && flush) cat = 9;
  else if (cnt[ord[0]] == 4) cat = 8;
  else if (cnt[ord[0]] == 3 && distinct == 2) cat = 7;
  else if (flush) cat = 6;
  else if (straight) cat = 5;
  else if (cnt[ord[0]] == 3) cat = 4;
  else if (cnt[ord[0]] == 2 && distinct == 3) cat = 3;
  else if (cnt[ord[0]] == 2) cat = 2;
  else cat = 1;
  uint32_t k = (uint32_t)cat << 20;
  if (straight) return k | ((uint32_t)top << 16);
  for (int i = 0; i < distinct; ++i) k |= (uint32_t)ord[i] << (16 - 4 * i);
  return k;
}

static uint32_t g_keys[7462];
static int g_nkeys = 0;
static int g_cat_start[11];

static int cmp_u32(const void* a, const void* b) {
  uint32_t x = *(const uint32_t*)a, y = *(const uint32_t*)b;
  return x < y ? -1 : x > y;
}

static void init_keys(void) {
  if (g_nkeys) return;
  uint32_t* all = (uint32_t*)malloc(2598960 * sizeof(uint32_t));
  int n = 0, c[5];
  for (c[0] = 1; c[0] <= 52; ++c[0])
    for (c[1] = c[0] + 1; c[1] <= 52; ++c[1])
      for (c[2] = c[1] + 1; c[2] <= 52; ++c[2])
        for (c[3] = c[2] + 1; c[3] <= 52; ++c[3])
          for (c[4] = c[3] + 1; c[4] <= 52; ++c[4]) all[n++] = key5(c);
  qsort(all, n, sizeof(uint32_t), cmp_u32);
  int m = 0;
  for (int i = 0; i < n; ++i)
    if (i == 0 || all[i] != all[i - 1]) g_keys[m++] = all[i];
  free(all);
  for (int cat = 0; cat <= 10; ++cat) g_cat_start[cat] = m;
  for (int i = m - 1; i >= 0; --i) g_cat_start[g_keys[i] >> 20] = i;
#pragma omp critical
  g_nkeys = m;
}

int oracle_num_hand_classes(void) { init_keys(); return g_nkeys; }

/* value of the best five of n (5..7) cards */
int oracle_eval_bruteforce(const int32_t* cards, int n) {
  init_keys();
  uint32_t best = 0;
  int c[5];
  for (int a = 0; a < n; ++a)
    for (int b = a + 1; b < n; ++b)
      for (int d = b + 1; d < n; ++d)
        for (int e = d + 1; e < n; ++e)
          for (int f = e + 1; f < n; ++f) {
            c[0] = cards[a]; c[1] = cards[b]; c[2] = cards[d]; c[3] = cards[e]; c[4] = cards[f];
            uint32_t k = key5(c);
            if (k > best) best = k;
          }
  uint32_t* p = (uint32_t*)bsearch(&best, g_keys, g_nkeys, sizeof(uint32_t), cmp_u32);
  int idx = (int)(p - g_keys);
  int cat = (int)(best >> 20);
  return (cat << 12) + (idx - g_cat_start[cat] + 1);
}

/* the reference's walk: p = 53; p = HR[p + card]...; 5/6 cards take one more HR[p]
 * (PokerGPU.py:437-444, :497-500) */
int oracle_handranks_walk(const int32_t* hr, const int32_t* cards, int n) {
  int p = 53;
  for (int i = 0; i < n; ++i) p = hr[p + cards[i]];
  if (n < 7) p = hr[p];
  return p;
}

/* Full check of a HandRanks table against the brute-force evaluator over every 7-card hand
 * (133,784,560; OpenMP) when stride == 1, or every stride-th hand.  compare != 0 also evaluates each
 * visited hand by brute force (21 five-card subsets).  Returns the number of mismatches; fills
 * cat_counts[10], min7, max7. */
long long oracle_check_table7(const int32_t* hr, int stride, int compare, long long* cat_counts, int* min7,
                              int* max7) {
  init_keys();
  long long bad = 0;
  long long cc[10] = {0};
  int mn = 1 << 30, mx = 0;
#pragma omp parallel for schedule(dynamic) reduction(+ : bad) reduction(min : mn) reduction(max : mx) reduction(+ : cc[:10])
  for (int a = 1; a <= 46; ++a) {
    long long idx = 0;
    int32_t c[7];
    c[0] = a;
    for (c[1] = a + 1; c[1] <= 47; ++c[1])
      for (c[2] = c[1] + 1; c[2] <= 48; ++c[2])
        for (c[3] = c[2] + 1; c[3] <= 49; ++c[3])
          for (c[4] = c[3] + 1; c[4] <= 50; ++c[4])
            for (c[5] = c[4] + 1; c[5] <= 51; ++c[5])
              for (c[6] = c[5] + 1; c[6] <= 52; ++c[6]) {
                if ((idx++ % stride) != 0) continue;
                int v = oracle_handranks_walk(hr, c, 7);
                cc[v >> 12 < 10 ? v >> 12 : 0]++;
                if (v < mn) mn = v;
                if (v > mx) mx = v;
                if (compare && v != oracle_eval_bruteforce(c, 7)) bad++;
              }
  }
  for (int i = 0; i < 10; ++i) cat_counts[i] = cc[i];
  *min7 = mn; *max7 = mx;
  return bad;
}

/* ---- Particle2D (environments/Particle2D/Particle2D.py:22-30) ------------------------------- */
void oracle_particle2d_step(float* state, const float* action, int32_t* steps, float* reward,
                            uint8_t* terminated, float dt, int max_steps, int n) {
#pragma omp parallel for if (n > 65536)
  for (int i = 0; i < n; ++i) {
    float ax = action[2 * i], ay = action[2 * i + 1];
    ax = ax < -1.f ? -1.f : (ax > 1.f ? 1.f : ax);                       /* :23 clamp */
    ay = ay < -1.f ? -1.f : (ay > 1.f ? 1.f : ay);
    float* s = state + 4 * i;
    volatile float t;
    t = ax * dt; s[2] = s[2] + t;                                         /* :24 vel += a*dt */
    t = ay * dt; s[3] = s[3] + t;
    t = s[2] * dt; s[0] = s[0] + t;                                       /* :25 pos += vel*dt */
    t = s[3] * dt; s[1] = s[1] + t;
    volatile float xx = s[0] * s[0], yy = s[1] * s[1];
    float dist = sqrtf(xx + yy);                                          /* :26 */
    volatile float a2x = ax * ax, a2y = ay * ay;
    volatile float pen = 0.001f * (a2x + a2y);
    reward[i] = -dist - pen;                                              /* :27 */
    steps[i] += 1;                                                        /* :28 */
    terminated[i] = (uint8_t)((dist < 0.1f) | (steps[i] >= max_steps));   /* :29 */
  }
}

/* ---- host threads used by the OpenMP loops (torchrun exports OMP_NUM_THREADS=1; the timed CPU baseline
 * sets the count explicitly and reports what it really got) ---------------------------------------- */
#ifdef _OPENMP
#include <omp.h>
int oracle_set_threads(int n) { if (n > 0) omp_set_num_threads(n); return omp_get_max_threads(); }
#else
int oracle_set_threads(int n) { (void)n; return 1; }
#endif

/* Enumerated hand strength -- no reference code computes it; this restates the DEFINITION the
 * engine's header gives (include/pulse_envs.h, pulse_poker_packed_enumerated_strength), built on the reference's river
 * quantity clamp((v7 - 4109) / 32765, 0, 1) (/root/reference/environments/Poker/PokerGPU.py:477-481): the mean of
 * that quantity over every completion of the board from the cards the seat cannot see.  Brute force with the
 * five-of-seven evaluator above; board = five cards (1..52), the first 3/4/5 visible at stage 1/2/3. */
float oracle_enumerated_strength(int stage, int h0, int h1, const int32_t* board) {
  if (stage < 1 || stage > 3) return 0.5f;
  int vis = stage + 2;
  int32_t c[7];
  char used[53] = {0};
  c[0] = h0; c[1] = h1; used[h0] = used[h1] = 1;
  for (int i = 0; i < vis; ++i) { c[2 + i] = board[i]; used[board[i]] = 1; }
  long long sum = 0, count = 0;
  if (stage == 3) {
    int v = oracle_eval_bruteforce(c, 7) - 4109;
    sum = v < 0 ? 0 : (v > 32765 ? 32765 : v); count = 1;
  } else if (stage == 2) {
    for (int x = 1; x <= 52; ++x) {
      if (used[x]) continue;
      c[6] = x;
      int v = oracle_eval_bruteforce(c, 7) - 4109;
      sum += v < 0 ? 0 : (v > 32765 ? 32765 : v); ++count;
    }
  } else {
    for (int x = 1; x <= 52; ++x) {
      if (used[x]) continue;
      for (int y = x + 1; y <= 52; ++y) {
        if (used[y]) continue;
        c[5] = x; c[6] = y;
        int v = oracle_eval_bruteforce(c, 7) - 4109;
        sum += v < 0 ? 0 : (v > 32765 ? 32765 : v); ++count;
      }
    }
  }
  return (float)sum / ((float)count * 32765.0f);
}

/* out [n, ap] from the reference-shaped arrays: stages [n], hands [n, P, 2], board [n, 5] (int32) */
void oracle_enumerated_strength_batch(const int32_t* stages, const int32_t* hands, const int32_t* board, int P, int ap,
                                      long long n, float* out) {
  init_keys();
#pragma omp parallel for schedule(dynamic, 16)
  for (long long g = 0; g < n; ++g)
    for (int s = 0; s < ap; ++s)
      out[g * ap + s] = oracle_enumerated_strength(stages[g], hands[(g * P + s) * 2], hands[(g * P + s) * 2 + 1],
                                                   board + g * 5);
}

/* Showdown equity (second half) -- no reference code computes it (its notes ask for equity "using ALL
 * players card info" and call it out of reach, /root/reference/rl/depth_notes.txt:316-321); this restates the DEFINITION
 * in include/pulse_envs.h (pulse_poker_packed_showdown_equity) from the reference-shaped arrays: for every completion of
 * the board from the cards in nobody's hand and not on the visible board, the contenders (status 0 ACTIVE or 2 ALLIN,
 * PokerGPU.py:11) compare seven-card values as the reference's showdown does (table walk, :437-444, when `hr` is given;
 * the brute-force evaluator above otherwise) and the best share one pot unit.  Flop / turn / river exact; preflop
 * `samples` boards, board q a forward Fisher-Yates over the unseen cards in (suit, rank) order driven by tape block
 * (seed, env_id, stream 8 | q << 4, counter): first two picks from word 0 (one draw in [0, m (m - 1))), then words 1..3. */
static int eq_value7(const int32_t* hr, const int32_t* c7) {
  return hr ? oracle_handranks_walk(hr, c7, 7) : oracle_eval_bruteforce(c7, 7);
}
static void eq_score(const int32_t* hr, const int32_t* hands, const int* contender, int ap, const int32_t* b5,
                     long long* num) {
  int val[10], best = -1, k = 0;
  for (int s = 0; s < ap; ++s) {
    if (!contender[s]) continue;
    int32_t c7[7] = {hands[2 * s], hands[2 * s + 1], b5[0], b5[1], b5[2], b5[3], b5[4]};
    val[s] = eq_value7(hr, c7);
    if (val[s] > best) best = val[s];
  }
  for (int s = 0; s < ap; ++s) if (contender[s] && val[s] == best) ++k;
  for (int s = 0; s < ap; ++s) if (contender[s] && val[s] == best) num[s] += 2520 / k;
}
static int eq_cmp_suit_rank(const void* a, const void* b) {
  int x = *(const int32_t*)a - 1, y = *(const int32_t*)b - 1;
  return ((x & 3) * 16 + (x >> 2)) - ((y & 3) * 16 + (y >> 2));
}
void oracle_showdown_equity(const int32_t* hr, const int32_t* hands, const int32_t* status, const int32_t* board, int ap,
                            int samples, uint64_t seed, uint64_t env_id, uint32_t counter, float* out) {
  char used[53] = {0};
  int contender[10], nc = 0;
  long long num[10] = {0}, boards = 0;
  for (int s = 0; s < ap; ++s) {
    used[hands[2 * s]] = used[hands[2 * s + 1]] = 1;
    contender[s] = status[s] == 0 || status[s] == 2;
    nc += contender[s];
  }
  int32_t b5[5];
  int vis = 0;
  for (int i = 0; i < 5; ++i) if (board[i] >= 1) { b5[vis++] = board[i]; used[board[i]] = 1; }
  int32_t unseen[52];
  int m = 0;
  for (int c = 1; c <= 52; ++c) if (!used[c]) unseen[m++] = c;
  qsort(unseen, m, sizeof(int32_t), eq_cmp_suit_rank);
  if (nc == 0) { for (int s = 0; s < ap; ++s) out[s] = 0.0f; return; }
  if (vis == 5) { eq_score(hr, hands, contender, ap, b5, num); boards = 1; }
  else if (vis == 4) {
    for (int i = 0; i < m; ++i) { b5[4] = unseen[i]; eq_score(hr, hands, contender, ap, b5, num); ++boards; }
  } else if (vis == 3) {
    for (int i = 0; i < m; ++i)
      for (int j = i + 1; j < m; ++j) { b5[3] = unseen[i]; b5[4] = unseen[j]; eq_score(hr, hands, contender, ap, b5, num); ++boards; }
  } else if (vis == 0) {
    for (int q = 0; q < samples; ++q) {
      uint32_t w[4];
      int32_t list[52];
      oracle_tape(seed, env_id, 8u | ((uint32_t)q << 4), counter, w);
      memcpy(list, unseen, sizeof(int32_t) * m);
      uint32_t a = (uint32_t)(((uint64_t)w[0] * (uint64_t)(m * (m - 1))) >> 32);
      int pick[5] = {(int)(a / (uint32_t)(m - 1)), 1 + (int)(a % (uint32_t)(m - 1)),
                     2 + (int)(((uint64_t)w[1] * (uint64_t)(m - 2)) >> 32), 3 + (int)(((uint64_t)w[2] * (uint64_t)(m - 3)) >> 32),
                     4 + (int)(((uint64_t)w[3] * (uint64_t)(m - 4)) >> 32)};
      for (int k = 0; k < 5; ++k) { int32_t x = list[pick[k]]; list[pick[k]] = list[k]; list[k] = x; b5[k] = x; }
      eq_score(hr, hands, contender, ap, b5, num);
      ++boards;
    }
  }
  for (int s = 0; s < ap; ++s) out[s] = boards ? (float)((double)num[s] / ((double)boards * 2520.0)) : 0.0f;
}

/* out [n, ap]; hands [n, P, 2], status [n, P], board [n, 5] (-1 = not dealt yet), reference-shaped int32 */
void oracle_showdown_equity_batch(const int32_t* hr, const int32_t* hands, const int32_t* status, const int32_t* board,
                                  int P, int ap, long long n, int samples, uint64_t seed, uint64_t env_base,
                                  uint32_t counter, float* out) {
  init_keys();
#pragma omp parallel for schedule(dynamic, 4)
  for (long long g = 0; g < n; ++g)
    oracle_showdown_equity(hr, hands + g * P * 2, status + g * P, board + g * 5, ap, samples, seed,
                           env_base + (uint64_t)g, counter, out + g * ap);
}

/* Exact preflop showdown equity of ONE table by full enumeration of C(m, 5) boards (heads-up: 1,712,304) -- the
 * yardstick the sampled preflop value is checked against.  Needs the HandRanks table. */
void oracle_showdown_equity_preflop_exact(const int32_t* hr, const int32_t* hands, const int32_t* status, int ap,
                                          double* out) {
  char used[53] = {0};
  int contender[10];
  for (int s = 0; s < ap; ++s) {
    used[hands[2 * s]] = used[hands[2 * s + 1]] = 1;
    contender[s] = status[s] == 0 || status[s] == 2;
  }
  int32_t u[52];
  int m = 0;
  for (int c = 1; c <= 52; ++c) if (!used[c]) u[m++] = c;
  long long tot[10] = {0}, boards = 0;
#pragma omp parallel
  {
    long long num[10] = {0}, nb = 0;
#pragma omp for schedule(dynamic, 1)
    for (int a = 0; a < m; ++a)
      for (int b = a + 1; b < m; ++b)
        for (int c = b + 1; c < m; ++c)
          for (int d = c + 1; d < m; ++d)
            for (int e = d + 1; e < m; ++e) {
              int32_t b5[5] = {u[a], u[b], u[c], u[d], u[e]};
              eq_score(hr, hands, contender, ap, b5, num);
              ++nb;
            }
#pragma omp critical
    { for (int s = 0; s < ap; ++s) tot[s] += num[s]; boards += nb; }
  }
  for (int s = 0; s < ap; ++s) out[s] = (double)tot[s] / ((double)boards * 2520.0);
}
