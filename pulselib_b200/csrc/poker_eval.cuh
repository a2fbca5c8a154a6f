// Computed 5/6/7-card hold'em evaluator producing the same values the TwoPlusTwo HandRanks table
// stores: value = (category << 12) + rank_in_category, category 1 high card .. 9 straight flush,
// rank ascending with strength (category sizes 1277, 2860, 858, 858, 10, 1277, 156, 156, 10).
// The reference walks a 130 MB table with 7 dependent gathers per seat
// (/root/reference/environments/Poker/PokerGPU.py:437-444, :477-478); on B200 the same integers come from
// four 13-bit suit masks in registers and at most ONE read of a small table that stays in L1:
//   all ranks distinct, no flush (high card / straight): pe_tab_unique[rank mask]            16 KB
//   flush / straight flush:                              pe_tab_flush[the suit's rank mask]  16 KB
//   one pair:                                            pe_tab_kick3[kickers, pair squeezed out]  8 KB
//   two pair, trips, full house, quads: a handful of integer operations, no table.
// The tables (tools/gen_eval_tables.py -> poker_eval_tables.inc) replaced the combinatorial-number-system arithmetic the
// first version did in registers: a warp's 32 hands fall into five or six categories, every category's code runs for
// the whole warp, and the binomials of the high-card and pair branches alone were a third of the deal kernel's
// instructions (profiles/r2t_reset_raw.csv: the evaluator was 59% of the whole-batch reset).  The table
// generator (handranks_gen.cpp) uses this very function for its leaves, and the tests check it
// against an independent brute-force evaluator in oracle/.
//
// Card ids are the table's: 1..52, rank-major, card = 4*rank + suit + 1, rank 0 = deuce .. 12 = ace.
#pragma once
#include <stdint.h>
#ifndef __CUDACC__
#define __host__
#define __device__
#define __forceinline__ inline
#endif

// one copy for host code and, under nvcc, one in device memory (read through the non-coherent path: L1-resident)
#define PE_TAB(name, n) static const uint16_t name##_h[n]
#include "poker_eval_tables.inc"
#undef PE_TAB
#ifdef __CUDACC__
#define PE_TAB(name, n) static __device__ const uint16_t name##_d[n]
#include "poker_eval_tables.inc"
#undef PE_TAB
#endif
#ifdef __CUDA_ARCH__
#define PE_LOOKUP(name, i) ((int)__ldg(&name##_d[i]))
#else
#define PE_LOOKUP(name, i) ((int)name##_h[i])
#endif

__host__ __device__ __forceinline__ int pe_hb(uint32_t x) {  // index of highest set bit, x != 0
#ifdef __CUDA_ARCH__
  return 31 - __clz((int)x);
#else
  return 31 - __builtin_clz(x);
#endif
}
__host__ __device__ __forceinline__ int pe_popc(uint32_t x) {
#ifdef __CUDA_ARCH__
  return __popc(x);
#else
  return __builtin_popcount(x);
#endif
}
__host__ __device__ __forceinline__ int pe_c2(int b) { return (b * (b - 1)) >> 1; }

// Straight rank 1..10 (wheel = 1, broadway = 10) of a 13-bit rank mask, or 0.
__host__ __device__ __forceinline__ int pe_straight(uint32_t m) {
  uint32_t mm = (m << 1) | ((m >> 12) & 1u);
  uint32_t run = mm & (mm >> 1) & (mm >> 2) & (mm >> 3) & (mm >> 4);
  return run ? pe_hb(run) + 1 : 0;
}

// any/pairs/trips/quads: rank masks with count >= 1 / == 2 / == 3 / == 4; flush: rank mask of the
// suit holding >= 5 cards, else 0.
__host__ __device__ __forceinline__ int pe_eval_core(uint32_t any, uint32_t pairs, uint32_t trips,
                                                     uint32_t quads, uint32_t flush) {
  if (flush) return PE_LOOKUP(pe_tab_flush, flush & 0x1FFFu);               // straight flush (9) or flush (6)
  if (!(pairs | trips | quads)) return PE_LOOKUP(pe_tab_unique, any & 0x1FFFu);  // straight (5) or high card (1)
  if (quads) {
    int q = pe_hb(quads);
    int k = pe_hb(any & ~(1u << q));
    return (8 << 12) + q * 12 + (k - (k > q)) + 1;
  }
  if (trips && ((trips & (trips - 1)) || pairs)) {
    int t = pe_hb(trips);
    int p = pe_hb((trips & ~(1u << t)) | pairs);
    return (7 << 12) + t * 12 + (p - (p > t)) + 1;
  }
  int st = pe_straight(any);            // six or seven cards: a straight next to a pair / two pair / trips
  if (st) return (5 << 12) + st;
  if (trips) {
    int t = pe_hb(trips);
    uint32_t s = any & ~(1u << t);
    int k1 = pe_hb(s); s ^= 1u << k1;
    int k2 = pe_hb(s);
    k1 -= (k1 > t); k2 -= (k2 > t);
    return (4 << 12) + t * 66 + pe_c2(k1) + k2 + 1;
  }
  if (pairs & (pairs - 1)) {
    int p1 = pe_hb(pairs);
    int p2 = pe_hb(pairs & ~(1u << p1));
    int k = pe_hb(any & ~(1u << p1) & ~(1u << p2));
    k -= (k > p1) + (k > p2);
    return (3 << 12) + (pe_c2(p1) + p2) * 11 + k + 1;
  }
  // one pair: the three best kickers, numbered as if the pair's rank did not exist
  int p = pe_hb(pairs);
  uint32_t s = any & ~(1u << p);
  s = (s & ((1u << p) - 1u)) | ((s >> (p + 1)) << p);
  return (2 << 12) + p * 220 + PE_LOOKUP(pe_tab_kick3, s & 0xFFFu) + 1;
}

// Evaluates 5..7 distinct cards given as four per-suit rank masks.
__host__ __device__ __forceinline__ int pe_eval_suits(uint32_t m0, uint32_t m1, uint32_t m2, uint32_t m3) {
  // bit-sliced per-rank count of the four suit bits
  uint32_t a = m0 & m1, b = m0 ^ m1, c = m2 & m3, d = m2 ^ m3;
  uint32_t ones = b ^ d, carry = b & d;
  uint32_t quads = a & c;
  uint32_t twos = (a ^ c) | carry;  // count in {2,3} (quads excluded: a^c == 0 and carry == 0 there)
  uint32_t any = m0 | m1 | m2 | m3;
  uint32_t trips = twos & ones, pairs = twos & ~ones;
  uint32_t flush = 0;
  if (pe_popc(m0) >= 5) flush = m0;
  if (pe_popc(m1) >= 5) flush = m1;
  if (pe_popc(m2) >= 5) flush = m2;
  if (pe_popc(m3) >= 5) flush = m3;
  return pe_eval_core(any, pairs, trips, quads, flush);
}

// Adds card (1..52) to the suit masks; ids outside 1..52 are ignored.
__host__ __device__ __forceinline__ void pe_add_card(int card, uint32_t m[4]) {
  unsigned c = (unsigned)(card - 1);
  if (c < 52u) m[c & 3u] |= 1u << (c >> 2);
}

// Same evaluation from one 64-bit word holding the four suit masks in 16-bit lanes
// (bit 16*suit + rank): adding a card is one shift and one OR.
__host__ __device__ __forceinline__ uint64_t pe_add_card64(int card, uint64_t m) {
  unsigned c = (unsigned)(card - 1);
  if (c < 52u) m |= 1ull << (((c & 3u) << 4) | (c >> 2));
  return m;
}
__host__ __device__ __forceinline__ int pe_eval_mask64(uint64_t m) {
  return pe_eval_suits((uint32_t)m & 0xFFFFu, (uint32_t)(m >> 16) & 0xFFFFu, (uint32_t)(m >> 32) & 0xFFFFu,
                       (uint32_t)(m >> 48));
}
