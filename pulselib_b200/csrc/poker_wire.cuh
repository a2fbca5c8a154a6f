// Compact per-table step record ("wire row") of the PokerGPU host-buffer path.
//
// The reference's step hands back a float32 observation row of W = 13 + 3 (M - 1) numbers per table plus reward and
// done (/root/reference/environments/Poker/PokerGPU.py:159-186, :633): 69 B per table-step heads-up.  With the caller's
// buffers in HOST memory that is what crosses PCIe, and it capped the end-to-end rate of round 1 at the link
// (SCALE_r01: 1.3e9 table-steps/s on eight GPUs).  Every number in the row is a small integer, and a table that was
// already finished when the step began repeats its previous row with reward 0.  So the device ships, per table that
// was LIVE when the step began, one packed record, compacted in table order; the host expander (csrc/host_expand.cpp)
// rebuilds the exact float32 row in the caller's buffer.
//
// The record is laid out for the HOST: the first eight observation columns as signed bytes, the rest as unsigned
// 16-bit words in observation order, so that expanding is "widen and convert" (a sign-extending and a zero-extending
// vector load plus two int->float converts for a heads-up row).  The first format of this round packed fields at
// 6- and 14-bit granularity into 24 B and cost 12 ns of scalar shifting per row on the host -- which bounded the
// whole path (profiles/r2c_bench_lines.jsonl: 20 ms of thread time per 4M-table step).
//
//   bytes 0-7   int8:  board[0..4], the actor's two hole cards (-1 = not dealt), stage                  = obs[0..7]
//   then u16:   (idx - button) mod AP, pot, highest - crb[idx], stack[idx], status[idx] | done << 15,
//               and per opponent k = 1 .. M-1 (seat (idx + k) mod AP; zeros beyond AP): stack, status, crb  = obs[8..W-1]
//               (padded to a whole 32-bit word)
//   last 4 B:   reward (float32 bits)
//   28 B heads-up, 40 B four seats, 52 B six seats, 76 B ten seats.
//
// Representable when every chip count fits 16 bits: stacks never exceed the table's total, n_players * max_bbs
// (busted / oversized stacks are only refilled to starting_bbs at reset, PokerGPU.py:104-108), which must be < 65536
// (the default 10 x 1000 is).  pulse_poker_host_create refuses other configurations.
#pragma once
#include "poker_core.cuh"

// 32-bit words per record
PULSE_HD constexpr int pkw_row_words(int M) { return 3 + (13 + 3 * (M - 1) - 8 + 1) / 2; }

// MT = max_players at compile time (row layout and every index below are then constants; the row stays in registers)
template <int MAXP, int MT>
PULSE_HD void pkw_pack_row(const PokerTable<MAXP>& t, const PokerParams& pr, float reward, uint32_t (&w)[pkw_row_words(MT)]) {
  constexpr int W = 13 + 3 * (MT - 1), T = W - 8, R = pkw_row_words(MT);
  const int AP = pr.AP;
  const int idx = t.idx;
  const int h0 = pk_sel(t.hole0, idx), h1 = pk_sel(t.hole1, idx);
  w[0] = (uint32_t)(t.board[0] & 0xFF) | (uint32_t)(t.board[1] & 0xFF) << 8 | (uint32_t)(t.board[2] & 0xFF) << 16 |
         (uint32_t)(t.board[3] & 0xFF) << 24;
  w[1] = (uint32_t)(t.board[4] & 0xFF) | (uint32_t)(h0 & 0xFF) << 8 | (uint32_t)(h1 & 0xFF) << 16 | (uint32_t)(t.stage & 0xFF) << 24;
  int rel = idx - t.button;
  if (rel < 0) rel += AP;
  if (rel >= AP) rel %= AP;
  uint32_t tail[T + 1];   // the u16 section, one value per observation column 8 .. W-1 (+ one pad)
  tail[0] = (uint32_t)rel;
  tail[1] = (uint32_t)(uint16_t)t.pot;
  tail[2] = (uint32_t)(uint16_t)(t.highest - pk_sel(t.crb, idx));
  tail[3] = (uint32_t)(uint16_t)pk_sel(t.stack, idx);
  tail[4] = (uint32_t)pk_status(t.st, idx) | (t.done ? 0x8000u : 0u);
  int seat = idx;
#pragma unroll
  for (int k = 1; k < MT; ++k) {
    uint32_t a = 0, b = 0, c = 0;
    if (k < AP) {
      seat = seat + 1 == AP ? 0 : seat + 1;
      a = (uint32_t)(uint16_t)pk_sel(t.stack, seat); b = (uint32_t)pk_status(t.st, seat); c = (uint32_t)(uint16_t)pk_sel(t.crb, seat);
    }
    tail[5 + 3 * (k - 1)] = a; tail[6 + 3 * (k - 1)] = b; tail[7 + 3 * (k - 1)] = c;
  }
  tail[T] = 0;
#pragma unroll
  for (int p = 0; p < T; p += 2) w[2 + p / 2] = tail[p] | tail[p + 1] << 16;
  union { float f; uint32_t u; } rb;
  rb.f = reward;
  w[R - 1] = rb.u;   // the row's last four bytes
}

// Portable host expansion of one record -> the reference's float32 row, reward, done (PokerGPU.py:159-179).
// csrc/host_expand.cpp holds the vectorised forms the host env uses; this one is the definition they are tested against.
inline void pkw_expand_row(const uint32_t* w, int M, float* o, float* reward, uint8_t* done) {
  const int W = 13 + 3 * (M - 1), R = pkw_row_words(M);
  const int8_t* b = reinterpret_cast<const int8_t*>(w);
  for (int i = 0; i < 8; ++i) o[i] = (float)b[i];
  const uint16_t* u = reinterpret_cast<const uint16_t*>(w + 2);
  for (int p = 0; p < W - 8; ++p) o[8 + p] = (float)u[p];
  o[12] = (float)(u[4] & 0x7FFFu);
  *done = (uint8_t)(u[4] >> 15);
  union { float f; uint32_t x; } rb;
  rb.x = w[R - 1];
  *reward = rb.f;
}
