// Host-side expansion of the packed step records (poker_wire.cuh) into the caller's float32 arrays.
#pragma once
#include <stdint.h>

struct PulseExpandArgs {
  const uint32_t* rows;     // the chunk's records (compacted in table order, or one per table when off == nullptr)
  const uint32_t* off;      // the chunk's per-32-table record offsets, or nullptr: slot = table - chunk_first
  uint32_t* live;           // whole-batch live mask, bit b of live[w] = table 32 w + b is not done; read, then rewritten
  uint32_t* prev_live;      // rewritten with the live mask as THIS call found it (= which tables have a record in this
                            // call's output; pulse_poker_host_live_mask)
  long long chunk_first, g0, g1;   // tables [g0, g1) of the chunk starting at chunk_first (g0 a multiple of 32)
  int row_words, obs_width, max_players;
  float* obs; float* reward; uint8_t* done;
  bool fresh;               // after a deal: every table has a record
  bool records_only;        // the caller keeps the packed records: only done flags and the live mask are maintained
};

// Expands the records of tables [g0, g1).  Picks an AVX2 body when the CPU has it (checked once), else portable code.
void pulse_expand_range(const PulseExpandArgs& a);
const char* pulse_expand_isa(void);

// 2048 host-buffer path (csrc/tfe_host_env.cu): packed boards (tile exponents, 4 bits per cell, cell (r, c) at bits
// 16 r + 4 c) -> the reference's int32 tile values [side, side] per board (TFE.py:121), boards [g0, g1).
void pulse_tfe_expand_range(const uint64_t* boards, int32_t* obs, long long g0, long long g1, int side);
