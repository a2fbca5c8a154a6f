// 2048 / TFE: per-board bodies of the batched step kernels.  Rules follow
// /root/reference/environments/2048/TFE.py (numba_rotate_with_buffer :38-44, squash_row_optimized
// :75-108, add_tile_numba :17-34, is_game_over_numba :48-67, TFE.reset :143-149, TFE.step :152-189);
// line cites are to that file.  The reference steps one int32 board of tile VALUES per object; here a
// board is one uint64 of tile EXPONENTS, 4 bits per cell, cell (r, c) at bits 16 r + 4 c (a 4-nibble
// row frame for every side 2..4; cells outside the side x side square stay 0).  Representable while
// every tile <= 2^15 = 32768: two 32768 tiles are left unmerged (outside the packed domain; the
// reference would make 65536).
//
// The reference rotates the board k = action times counter-clockwise, squashes every row to the left
// and rotates back (:154-178).  That is: action 0 = left, 1 = up, 2 = right, 3 = down, done here as
// optional transpose / row reversal around one left-squash of the four 16-bit rows.  A tile is
// spawned after EVERY step, even when nothing moved (:181-182; changed_buffer is computed and unused).
#pragma once
#include "common.cuh"

#define TFE_FOUR_THRESHOLD 3865470566u  // floor(0.9 * 2^32): u1 > this  <=>  "random.random() > 0.9" (:33)

// 4x4 nibble transpose of a row-major board.  Device: two delta swaps inside the 32-bit halves (rows
// 0-1, rows 2-3) transpose the 2x2 nibble blocks, then two byte permutes (PRMT) exchange the
// off-diagonal 2x2 blocks between the halves -- 10 ALU operations instead of ~20 64-bit ones.
PULSE_HD uint64_t tfe_transpose(uint64_t x) {
#ifdef __CUDA_ARCH__
  uint32_t lo = (uint32_t)x, hi = (uint32_t)(x >> 32);
  uint32_t t = (lo ^ (lo >> 12)) & 0x0000F0F0u; lo ^= t ^ (t << 12);
  t = (hi ^ (hi >> 12)) & 0x0000F0F0u; hi ^= t ^ (t << 12);
  const uint32_t nlo = __byte_perm(lo, hi, 0x6240), nhi = __byte_perm(lo, hi, 0x7351);
  return ((uint64_t)nhi << 32) | nlo;
#else
  uint64_t a1 = x & 0xF0F00F0FF0F00F0FULL, a2 = x & 0x0000F0F00000F0F0ULL, a3 = x & 0x0F0F00000F0F0000ULL;
  uint64_t a = a1 | (a2 << 12) | (a3 >> 12);
  uint64_t b1 = a & 0xFF00FF0000FF00FFULL, b2 = a & 0x00FF00FF00000000ULL, b3 = a & 0x00000000FF00FF00ULL;
  return b1 | (b2 >> 24) | (b3 << 24);
#endif
}

// reverse the SIDE cells of every row
template <int SIDE>
PULSE_HD uint64_t tfe_reverse_rows(uint64_t x) {
  if (SIDE == 4) {
#ifdef __CUDA_ARCH__
    uint32_t lo = (uint32_t)x, hi = (uint32_t)(x >> 32);   // swap nibbles in bytes, then bytes in rows (PRMT)
    lo = ((lo << 4) & 0xF0F0F0F0u) | ((lo >> 4) & 0x0F0F0F0Fu);
    hi = ((hi << 4) & 0xF0F0F0F0u) | ((hi >> 4) & 0x0F0F0F0Fu);
    return ((uint64_t)__byte_perm(hi, 0, 0x2301) << 32) | __byte_perm(lo, 0, 0x2301);
#else
    x = ((x & 0x0F0F0F0F0F0F0F0FULL) << 4) | ((x >> 4) & 0x0F0F0F0F0F0F0F0FULL);
    return ((x & 0x00FF00FF00FF00FFULL) << 8) | ((x >> 8) & 0x00FF00FF00FF00FFULL);
#endif
  } else if (SIDE == 3) {
    return (x & 0x00F000F000F000F0ULL) | ((x & 0x000F000F000F000FULL) << 8) | ((x >> 8) & 0x000F000F000F000FULL);
  } else {
    return ((x & 0x000F000F000F000FULL) << 4) | ((x >> 4) & 0x000F000F000F000FULL);
  }
}

// squash_row_optimized (:75-108) on one 16-bit row of exponents: slide left, merge equal neighbours
// once per pair; score += merged tile value
PULSE_HD uint32_t tfe_squash_row(uint32_t row, uint32_t& score) {
  uint32_t a0 = row & 15u, a1 = (row >> 4) & 15u, a2 = (row >> 8) & 15u, a3 = (row >> 12) & 15u;
  if (!a2) { a2 = a3; a3 = 0; }
  if (!a1) { a1 = a2; a2 = a3; a3 = 0; }
  if (!a0) { a0 = a1; a1 = a2; a2 = a3; a3 = 0; }
  if (a0 && a0 == a1 && a0 < 15u) { a0 += 1; score += 1u << a0; a1 = a2; a2 = a3; a3 = 0; }
  if (a1 && a1 == a2 && a1 < 15u) { a1 += 1; score += 1u << a1; a2 = a3; a3 = 0; }
  if (a2 && a2 == a3 && a2 < 15u) { a2 += 1; score += 1u << a2; a3 = 0; }
  return a0 | (a1 << 4) | (a2 << 8) | (a3 << 12);
}

template <int SIDE>
PULSE_HD uint64_t tfe_cell_mask() {  // 0x1 in every real cell's nibble
  const uint64_t row = SIDE == 4 ? 0x1111ULL : SIDE == 3 ? 0x0111ULL : 0x0011ULL;
  uint64_t m = 0;
  for (int r = 0; r < SIDE; ++r) m |= row << (16 * r);
  return m;
}

// 0x1 in the nibble of every cell whose 4 bits are all zero (within `cells`)
PULSE_HD uint64_t tfe_zero_nibbles(uint64_t x, uint64_t cells) {
  x |= x >> 1; x |= x >> 2;
  return ~x & cells;
}

// add_tile_numba (:17-34) with the two RNG draws replaced by two 32-bit words:
// idx = mulhi32(u0, n_empty) among the empty cells in row-major order, four = u1 > TFE_FOUR_THRESHOLD
template <int SIDE>
PULSE_HD uint64_t tfe_add_tile(uint64_t b, uint32_t u0, uint32_t u1) {
  const uint64_t e = tfe_zero_nibbles(b, tfe_cell_mask<SIDE>());
  const uint32_t lo = (uint32_t)e, hi = (uint32_t)(e >> 32);
  const int nlo = PULSE_POPC(lo), ne = nlo + PULSE_POPC(hi);
  if (ne == 0) return b;
  int idx = (int)pulse_mulhi32(u0, (uint32_t)ne);
  // position of the idx-th set bit (set bits sit at multiples of 4): halve the window four times
  int pos = 0;
  uint32_t w = lo;
  if (idx >= nlo) { idx -= nlo; w = hi; pos = 32; }
  int c = PULSE_POPC(w & 0xFFFFu);
  if (idx >= c) { idx -= c; w >>= 16; pos += 16; }
  c = PULSE_POPC(w & 0xFFu);
  if (idx >= c) { idx -= c; w >>= 8; pos += 8; }
  c = (int)(w & 1u);
  if (idx >= c) pos += 4;
  return b | ((uint64_t)(u1 > TFE_FOUR_THRESHOLD ? 2u : 1u) << pos);  // exponent 2 (tile 4) or 1 (tile 2)
}

// is_game_over_numba (:48-67)
template <int SIDE>
PULSE_HD bool tfe_game_over(uint64_t b) {
  const uint64_t cells = tfe_cell_mask<SIDE>();
  if (tfe_zero_nibbles(b, cells)) return false;                         // an empty cell
  const uint64_t hpairs = cells & (cells >> 4) & 0x0FFF0FFF0FFF0FFFULL;  // cells with a right neighbour in the same row
  if (tfe_zero_nibbles(b ^ (b >> 4), hpairs)) return false;             // equal horizontal neighbours
  const uint64_t vpairs = cells & (cells >> 16);                        // cells with a neighbour below
  if (tfe_zero_nibbles(b ^ (b >> 16), vpairs)) return false;            // equal vertical neighbours
  return true;
}

// Where a row lives in the staged table.  Tile exponents are small (most cells hold 0..3), so the raw row
// value puts nearly every lookup of a warp into the same ~8 shared-memory banks with different addresses
// (6.9 wavefronts per lookup measured, profiles/r1q_tfe_step_raw.csv).  Folding bits 6.. and 11.. into the
// low bits (a bijection on 16 bits: bit i of the slot = row[i] ^ row[i+6] ^ row[i+11]) spreads the same
// lookups over all 32 banks (3.1-3.4 wavefronts on boards from random play).
PULSE_HD uint32_t tfe_row_slot(uint32_t row) { return (row ^ (row >> 6) ^ (row >> 11)) & 0xFFFFu; }
// tfe_row_slot on the four 16-bit rows of a board at once
PULSE_HD uint64_t tfe_board_slots(uint64_t b) {
#ifdef __CUDA_ARCH__
  uint32_t lo = (uint32_t)b, hi = (uint32_t)(b >> 32);
  lo = lo ^ ((lo >> 6) & 0x03FF03FFu) ^ ((lo >> 11) & 0x001F001Fu);
  hi = hi ^ ((hi >> 6) & 0x03FF03FFu) ^ ((hi >> 11) & 0x001F001Fu);
  return ((uint64_t)hi << 32) | lo;
#else
  return b ^ ((b >> 6) & 0x03FF03FF03FF03FFULL) ^ ((b >> 11) & 0x001F001F001F001FULL);
#endif
}

// How a 16-bit row is squashed: computed (host and device), or looked up in the 65,536-entry row-move
// table the step kernel stages in shared memory (tfe.cu).  Both give squash_row_optimized's result; the
// table is built from tfe_squash_row.  slots() maps a board's rows to what at() is indexed by.
struct TfeComputedRows {
  PULSE_HD uint64_t slots(uint64_t b) const { return b; }
  PULSE_HD uint32_t at(uint32_t row, uint32_t& score) const { return tfe_squash_row(row, score); }
  PULSE_HD uint32_t operator()(uint32_t row, uint32_t& score) const { return tfe_squash_row(row, score); }
};
// score code of a row: the exponents of its (at most two) merged tiles, 4 bits each, 0 = none
PULSE_HD uint32_t tfe_score_code(uint32_t row) {
  uint32_t a0 = row & 15u, a1 = (row >> 4) & 15u, a2 = (row >> 8) & 15u, a3 = (row >> 12) & 15u;
  if (!a2) { a2 = a3; a3 = 0; }
  if (!a1) { a1 = a2; a2 = a3; a3 = 0; }
  if (!a0) { a0 = a1; a1 = a2; a2 = a3; a3 = 0; }
  uint32_t code = 0;
  if (a0 && a0 == a1 && a0 < 15u) { code = a0 + 1; a1 = a2; a2 = a3; a3 = 0; }
  if (a1 && a1 == a2 && a1 < 15u) { code = code ? code | ((a1 + 1) << 4) : a1 + 1; a2 = a3; a3 = 0; }
  if (a2 && a2 == a3 && a2 < 15u) { code = code ? code | ((a2 + 1) << 4) : a2 + 1; }
  return code;
}
// the score a code stands for: exponent 0 means "no merge" and a merged tile is at least 4, so the stray 1 << 0
// terms (at most 1 + 1) fall in the two low bits
PULSE_HD uint32_t tfe_code_score(uint32_t c) { return ((1u << (c & 15u)) + (1u << (c >> 4))) & ~3u; }
struct TfeTableRows {
  const uint16_t* result;        // [65536] squashed row, at tfe_row_slot(row)
  const uint8_t* code;           // [65536] tfe_score_code, at tfe_row_slot(row)
  const uint32_t* score_of_code; // [256] tfe_code_score: one more (mostly broadcast) read instead of six ALU operations per row
  PULSE_HD uint64_t slots(uint64_t b) const { return tfe_board_slots(b); }
  PULSE_HD uint32_t at(uint32_t slot, uint32_t& score) const {
    score += score_of_code[code[slot]];
    return result[slot];
  }
  PULSE_HD uint32_t operator()(uint32_t row, uint32_t& score) const { return at(tfe_row_slot(row), score); }
};

// the move of TFE.step (:154-178); returns the new board, adds the step's score
template <int SIDE, class Rows>
PULSE_HD uint64_t tfe_move(uint64_t b, int action, uint32_t& score, const Rows& rows) {
  const bool vertical = action & 1, toward_end = (action & 2) != 0;     // 1 up, 3 down; 2 right, 3 down
  if (vertical) b = tfe_transpose(b);
  if (toward_end) b = tfe_reverse_rows<SIDE>(b);
  const uint64_t slots = rows.slots(b);
  uint64_t out = 0;
#pragma unroll
  for (int r = 0; r < SIDE; ++r)
    out |= (uint64_t)rows.at((uint32_t)(slots >> (16 * r)) & 0xFFFFu, score) << (16 * r);
  if (toward_end) out = tfe_reverse_rows<SIDE>(out);
  if (vertical) out = tfe_transpose(out);
  return out;
}

// TFE.reset (:143-149): empty board, two spawns; c = (u0,u1) of spawn 1, (u0,u1) of spawn 2
template <int SIDE>
PULSE_HD uint64_t tfe_reset_board(PulseU4 c) {
  return tfe_add_tile<SIDE>(tfe_add_tile<SIDE>(0, c.x, c.y), c.z, c.w);
}

// One step of one board (:152-189) with the batched-engine reset folded in.  reset_now (auto_reset and the
// board was game-over when the step began): the board is cleared first, so the move is a no-op scoring 0,
// the spawn every step makes (c.x, c.y) is TFE.reset's first tile and a second spawn (c.z, c.w) is its
// second (:143-149); that step returns reward 0, done 0.  One code path for both cases, so a warp holding
// a finished board does not run a second instruction stream.
template <int SIDE, class Rows>
PULSE_HD uint64_t tfe_step_fused(uint64_t b, int action, bool reset_now, PulseU4 c, int64_t& score_total,
                                 int32_t& reward, bool& done, const Rows& rows) {
  if (reset_now) { b = 0; score_total = 0; }
  uint32_t score = 0;
  b = tfe_move<SIDE>(b, action & 3, score, rows);
  b = tfe_add_tile<SIDE>(b, c.x, c.y);                                  // :182 always
  if (reset_now) b = tfe_add_tile<SIDE>(b, c.z, c.w);
  score_total += score;                                                 // :168
#ifdef __CUDA_ARCH__
  reward = score ? 31 - __clz((int)score) : 0;                          // int(step_score).bit_length() - 1 (:185-187)
#else
  reward = score ? 31 - __builtin_clz(score) : 0;
#endif
  done = tfe_game_over<SIDE>(b);                                        // :189
  return b;
}

// the same with the two chance blocks handed in separately (spawn: x, y; reset: x, y, z, w)
template <int SIDE, class Rows>
PULSE_HD uint64_t tfe_step_one(uint64_t b, int action, bool was_done, int auto_reset, PulseU4 spawn /*x,y*/,
                               PulseU4 reset_c, int64_t& score_total, int32_t& reward, bool& done, const Rows& rows) {
  const bool reset_now = auto_reset && was_done;
  return tfe_step_fused<SIDE>(b, action, reset_now, reset_now ? reset_c : spawn, score_total, reward, done, rows);
}

// packed exponents -> the reference's int32 tile values, row-major side x side
template <int SIDE>
PULSE_HD void tfe_unpack_board(uint64_t b, int32_t* out) {
#pragma unroll
  for (int r = 0; r < SIDE; ++r)
#pragma unroll
    for (int c = 0; c < SIDE; ++c) {
      const uint32_t e = (uint32_t)(b >> (16 * r + 4 * c)) & 15u;
      out[r * SIDE + c] = e ? (int32_t)(1u << e) : 0;
    }
}

// int32 tile values -> packed exponents (boards handed in from outside); returns false when a value
// is not 0 or a power of two in 2..32768
template <int SIDE>
PULSE_HD bool tfe_pack_board(const int32_t* in, uint64_t& b) {
  b = 0;
  bool ok = true;
  for (int r = 0; r < SIDE; ++r)
    for (int c = 0; c < SIDE; ++c) {
      const int32_t v = in[r * SIDE + c];
      uint32_t e = 0;
      if (v != 0) {
        while (e < 16 && (1 << e) != v) ++e;
        if (e < 1 || e > 15) { ok = false; e = 0; }
      }
      b |= (uint64_t)e << (16 * r + 4 * c);
    }
  return ok;
}
