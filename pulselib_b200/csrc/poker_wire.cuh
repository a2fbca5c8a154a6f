// Compact per-table step record ("wire row") of the PokerGPU host-buffer path.
//
// The reference's step hands back a float32 observation row of 13 + 3 (M - 1) numbers per table plus reward and done
// (/root/reference/environments/Poker/PokerGPU.py:159-186, :633): 69 B per table-step heads-up.  With the caller's
// buffers in HOST memory that is what crosses PCIe, and it capped the end-to-end rate of round 1 at the link
// (SCALE_r01: 1.3e9 table-steps/s on eight GPUs).  Every number in the row is a small integer, and a table that was
// already finished when the step began repeats its previous row with reward 0.  So the device ships, per table that
// was LIVE when the step began, one packed record, compacted in table order; the host expander (csrc/host_env.cu)
// rebuilds the exact float32 row in the caller's buffer.
//
//   word 0 (u64): bits 0-29  five board cards, 6 bits each, 0 = not dealt (-1)
//                 bits 30-41 the actor's two hole cards, 6 bits each, 0 = none (-1)
//                 bits 42-44 stage, 45-48 (idx - button) mod AP, 49-50 the actor's status, 51 done
//   then u16s:    pot, highest - crb[idx], stack[idx], and per opponent k = 1 .. AP-1 (seat (idx + k) mod AP):
//                 stack | status << 14,  crb
//   last 4 bytes: reward (float32 bits)
//   row = 18 + 4 (M - 1) bytes rounded up to 8: 24 B heads-up, 40 B six seats, 56 B ten seats.
//
// Representable when every chip count fits: stacks never exceed the table's total, n_players * max_bbs
// (busted / oversized stacks are only refilled to starting_bbs at reset, PokerGPU.py:104-108), which must be < 16384
// (the default 10 x 1000 is).  pulse_poker_host_create refuses other configurations.
#pragma once
#include "poker_core.cuh"

PULSE_HD int pkw_row_words(int M) { return (18 + 4 * (M - 1) + 7) / 8; }

// cards are 1..52 or -1; 0 encodes -1
PULSE_HD uint32_t pkw_card(int c) { return c > 0 ? (uint32_t)c : 0u; }

// MT = max_players at compile time (row layout and every index below are then constants; the row stays in registers)
template <int MAXP, int MT>
PULSE_HD void pkw_pack_row(const PokerTable<MAXP>& t, const PokerParams& pr, float reward, uint64_t (&w)[(18 + 4 * (MT - 1) + 7) / 8]) {
  constexpr int R = (18 + 4 * (MT - 1) + 7) / 8;
  const int AP = pr.AP;
  const int idx = t.idx;
  uint64_t h = 0;
#pragma unroll
  for (int i = 0; i < 5; ++i) h |= (uint64_t)pkw_card(t.board[i]) << (6 * i);
  h |= (uint64_t)pkw_card(pk_sel(t.hole0, idx)) << 30;
  h |= (uint64_t)pkw_card(pk_sel(t.hole1, idx)) << 36;
  int rel = idx - t.button;
  if (rel < 0) rel += AP;
  if (rel >= AP) rel %= AP;
  h |= (uint64_t)(t.stage & 7) << 42 | (uint64_t)(rel & 15) << 45 | (uint64_t)pk_status(t.st, idx) << 49 |
       (uint64_t)(t.done ? 1 : 0) << 51;
#pragma unroll
  for (int i = 1; i < R; ++i) w[i] = 0;
  w[0] = h;
  // u16 number p of the chip section lives in word 1 + p / 4, lane p % 4
  w[1] = (uint64_t)(uint16_t)t.pot | (uint64_t)(uint16_t)(t.highest - pk_sel(t.crb, idx)) << 16 |
         (uint64_t)(uint16_t)pk_sel(t.stack, idx) << 32;
  int seat = idx;
#pragma unroll
  for (int k = 1; k < MT; ++k) {
    uint32_t a = 0, c = 0;
    if (k < AP) {
      seat = seat + 1 == AP ? 0 : seat + 1;
      a = (uint32_t)(pk_sel(t.stack, seat) & 0x3FFF) | (uint32_t)pk_status(t.st, seat) << 14;
      c = (uint32_t)(uint16_t)pk_sel(t.crb, seat);
    }
    const int pa = 3 + 2 * (k - 1), pc = pa + 1;   // constants once the loop is unrolled
    w[1 + pa / 4] |= (uint64_t)a << (16 * (pa % 4));
    w[1 + pc / 4] |= (uint64_t)c << (16 * (pc % 4));
  }
  union { float f; uint32_t u; } rb;
  rb.f = reward;
  w[R - 1] |= (uint64_t)rb.u << 32;   // the row's last four bytes
}

// Host side: one record -> the reference's float32 row, reward, done (PokerGPU.py:159-179).
inline void pkw_expand_row(const uint64_t* w, int AP, int M, float* o, float* reward, uint8_t* done) {
  const int R = pkw_row_words(M);
  const uint64_t h = w[0];
  for (int i = 0; i < 5; ++i) { int c = (int)((h >> (6 * i)) & 63u); o[i] = c ? (float)c : -1.0f; }
  { int c = (int)((h >> 30) & 63u); o[5] = c ? (float)c : -1.0f; }
  { int c = (int)((h >> 36) & 63u); o[6] = c ? (float)c : -1.0f; }
  o[7] = (float)((h >> 42) & 7u);
  o[8] = (float)((h >> 45) & 15u);
  o[12] = (float)((h >> 49) & 3u);
  *done = (uint8_t)((h >> 51) & 1u);
  const uint16_t* u = reinterpret_cast<const uint16_t*>(w + 1);
  o[9] = (float)u[0];
  o[10] = (float)u[1];
  o[11] = (float)u[2];
  for (int k = 1; k < M; ++k) {
    float a = 0.0f, b = 0.0f, c = 0.0f;
    if (k < AP) {
      const uint16_t s = u[3 + 2 * (k - 1)];
      a = (float)(s & 0x3FFF); b = (float)(s >> 14); c = (float)u[4 + 2 * (k - 1)];
    }
    o[13 + 3 * (k - 1)] = a; o[14 + 3 * (k - 1)] = b; o[15 + 3 * (k - 1)] = c;
  }
  union { float f; uint32_t x; } rb;
  rb.x = (uint32_t)(w[R - 1] >> 32);
  *reward = rb.f;
}
