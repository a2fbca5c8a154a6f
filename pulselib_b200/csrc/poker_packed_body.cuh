// Packed PokerGPU layout: the throughput path.  Same per-table logic as the reference-shaped
// kernels (poker_core.cuh), different HBM layout:
//
//   tbl   uint4 [N]    x = meta  (stage:3 | idx:4 | agg:4 | button:4 | acted:5 | done:1 | eq_stage:2)
//                          eq_stage: the street (1..3) on which a step last began, i.e. the street the
//                          reference's cached equities belong to (PokerGPU.py:542, :604-616); 0 = none yet
//                      y = status bit-planes (bit s = low bit of seat s, bit 16+s = high bit)
//                      z = pot, w = highest
//   tbl2  uint2 [N]    x = last_raise_size, y = the hand's five board cards, 6 bits each
//   stack/crb/inv  int32 [P][N]   seat-major, so lane i reads element i of every seat row
//   hole  uint16 [P][N]           card0 | card1 << 6
//   ev    uint64 [P][N]           the seat's hand values for this deal, computed once by reset:
//                                 bits 0-15 seven-card value (river equity and showdown rank),
//                                 16-31 six-card value (turn), 32-52 HR[five-card value] (flop)
//   hand_ctr uint32 [N]           hands dealt to this table (tape counter for auto-reset)
//
// One thread owns one table: a 16 B + 8 B vector load plus 3 coalesced 4 B loads per seat bring the
// whole table into registers; only words that changed are stored.  The deck itself is never kept:
// hole cards and the full board are fixed by the deal (burn/flop/turn/river sit at deck positions
// 2AP+1..2AP+7 whatever the betting does, PokerGPU.py:394-419, :601-616), so reset stores the five
// board cards and `stage` says how many are visible.  Heads-up state is 52 B per table against the
// 62 B the SURVEY's accounting assumes.  Because the whole board is known at the deal, so is every
// hand value the hand can ever need: reset evaluates each seat's 5-, 6- and 7-card hands once
// (uniformly across the warp) and step only reads `ev` -- one 8 B word for the actor when a reward
// needs its equity, one per live seat at showdown.  This equals the reference's cached equities
// because it refreshes its cache whenever stage or board change (PokerGPU.py:604-616, :542), and it
// takes the evaluator (the most divergent code) out of the per-step kernel.
#pragma once
#include "poker_core.cuh"
#include "poker_policy.cuh"
#include "poker_wire.cuh"

#define PKP_REDEAL_MARK 0x80000000u  // top bit of hand_ctr: table waits for the redeal pass

struct PackedDeck {
  uint32_t fb;   // five board cards, 6 bits each
  int base;      // 2 * active_players: first deck position after the hole cards
  PULSE_HD int card(int pos) const {
    int off = pos - base;  // 1,2,3 flop; 5 turn; 7 river
    int i = off <= 3 ? off - 1 : (off == 5 ? 3 : 4);
    return (int)((fb >> (6 * i)) & 63u);
  }
};

PULSE_HD uint32_t pkp_meta(int stage, int idx, int agg, int button, int acted, bool done, int eq_stage) {
  if (acted > 31) acted = 31;  // only ever compared with the number of ACTIVE seats (<= 10)
  return (uint32_t)stage | ((uint32_t)idx << 3) | ((uint32_t)agg << 7) | ((uint32_t)button << 11) |
         ((uint32_t)acted << 15) | ((uint32_t)done << 20) | ((uint32_t)eq_stage << 21);
}

struct PulseU4x { uint32_t x, y, z, w; };
struct PulseU2x { uint32_t x, y; };

PULSE_HD PulseU4x pkp_ld4(const uint32_t* p) {
#ifdef __CUDA_ARCH__
  uint4 v = *reinterpret_cast<const uint4*>(p);
  PulseU4x r; r.x = v.x; r.y = v.y; r.z = v.z; r.w = v.w; return r;
#else
  PulseU4x r; r.x = p[0]; r.y = p[1]; r.z = p[2]; r.w = p[3]; return r;
#endif
}
PULSE_HD void pkp_st4(uint32_t* p, uint32_t x, uint32_t y, uint32_t z, uint32_t w) {
#ifdef __CUDA_ARCH__
  *reinterpret_cast<uint4*>(p) = make_uint4(x, y, z, w);
#else
  p[0] = x; p[1] = y; p[2] = z; p[3] = w;
#endif
}
PULSE_HD PulseU2x pkp_ld2(const uint32_t* p) {
#ifdef __CUDA_ARCH__
  uint2 v = *reinterpret_cast<const uint2*>(p);
  PulseU2x r; r.x = v.x; r.y = v.y; return r;
#else
  PulseU2x r; r.x = p[0]; r.y = p[1]; return r;
#endif
}
PULSE_HD void pkp_st2(uint32_t* p, uint32_t x, uint32_t y) {
#ifdef __CUDA_ARCH__
  *reinterpret_cast<uint2*>(p) = make_uint2(x, y);
#else
  p[0] = x; p[1] = y;
#endif
}

// Ordered loads of the step's first-level words.  ptxas schedules for the longest dependency chain (table word ->
// actor seat -> that seat's hand-value word), so with ordinary loads it put the first consumer of the table word
// ahead of most per-seat loads: the warp waited for the table word, then issued the seat loads and waited again --
// two memory round trips per tile instead of one (SASS of r1r: LOP3 on the tbl register before 16 of 24 seat LDGs).
// Volatile loads keep program order among themselves, and the table word is loaded LAST, so nothing can consume it
// before every other first-level load has been issued.  (These words are read once per launch: bypassing L1 costs nothing.)
#ifdef __CUDA_ARCH__
__device__ __forceinline__ int32_t pkp_ldv32(const int32_t* p) { return *reinterpret_cast<const volatile int32_t*>(p); }
__device__ __forceinline__ uint32_t pkp_ldv16(const uint16_t* p) { return *reinterpret_cast<const volatile uint16_t*>(p); }
__device__ __forceinline__ PulseU2x pkp_ldv2(const uint32_t* p) {
  PulseU2x r;
  asm volatile("ld.volatile.global.v2.u32 {%0, %1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p));
  return r;
}
__device__ __forceinline__ PulseU4x pkp_ldv4(const uint32_t* p) {
  PulseU4x r;
  asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}
#else
inline int32_t pkp_ldv32(const int32_t* p) { return *p; }
inline uint32_t pkp_ldv16(const uint16_t* p) { return *p; }
inline PulseU2x pkp_ldv2(const uint32_t* p) { return pkp_ld2(p); }
inline PulseU4x pkp_ldv4(const uint32_t* p) { return pkp_ld4(p); }
#endif

template <int MAXP>
PULSE_HD void packed_load(PokerTable<MAXP>& t, uint32_t& fb, const PulsePokerPacked& S, long long g, long long n,
                          int AP) {
  PulseU4x a; PulseU2x b;
  uint32_t h[MAXP];
  if (MAXP <= 6) {
    // seats beyond AP re-read seat 0 and are zeroed by a select: no branch per seat, the loads form one straight line
    int32_t st[MAXP], cr[MAXP], iv[MAXP];
#pragma unroll
    for (int s = 0; s < MAXP; ++s) {
      const long long o = (long long)(s < AP ? s : 0) * n + g;
      st[s] = pkp_ldv32(S.stack + o); cr[s] = pkp_ldv32(S.crb + o); iv[s] = pkp_ldv32(S.inv + o);
      h[s] = pkp_ldv16(S.hole + o);
    }
    b = pkp_ldv2(S.tbl2 + 2 * g);
    a = pkp_ldv4(S.tbl + 4 * g);
#pragma unroll
    for (int s = 0; s < MAXP; ++s) {
      const bool on = s < AP;
      t.stack[s] = on ? st[s] : 0; t.crb[s] = on ? cr[s] : 0; t.inv[s] = on ? iv[s] : 0; h[s] = on ? h[s] : 0xFFFFFFFFu;
    }
  } else {
    // wider tables: holding every seat's words live at once costs more in spills than the strict order gains
    // (ten seats: 0.80 ordered against 0.84 like this) -- ordinary loads, written ahead of the decode
    a = pkp_ld4(S.tbl + 4 * g);
    b = pkp_ld2(S.tbl2 + 2 * g);
#pragma unroll
    for (int s = 0; s < MAXP; ++s) {
      t.stack[s] = t.crb[s] = t.inv[s] = 0; h[s] = 0xFFFFFFFFu;
      if (s < AP) {
        long long o = (long long)s * n + g;
        t.stack[s] = S.stack[o]; t.crb[s] = S.crb[o]; t.inv[s] = S.inv[o];
        h[s] = S.hole[o];
      }
    }
  }
#pragma unroll
  for (int s = 0; s < MAXP; ++s) {  // a seat that is not dealt in holds (-1, -1)
    t.hole0[s] = h[s] == 0xFFFFFFFFu ? -1 : (int)(h[s] & 63u);
    t.hole1[s] = h[s] == 0xFFFFFFFFu ? -1 : (int)((h[s] >> 6) & 63u);
  }
  t.stage = a.x & 7; t.idx = (a.x >> 3) & 15; t.agg = (a.x >> 7) & 15; t.button = (a.x >> 11) & 15;
  t.acted = (a.x >> 15) & 31; t.done = (a.x >> 20) & 1; t.eq_stage = (a.x >> 21) & 3; t.dirty = false;
  t.st = a.y; t.pot = (int32_t)a.z; t.highest = (int32_t)a.w; t.last_raise = (int32_t)b.x;
  fb = b.y;
  const int vis = t.stage == 0 ? 0 : t.stage == 1 ? 3 : t.stage == 2 ? 4 : 5;
#pragma unroll
  for (int i = 0; i < 5; ++i) t.board[i] = i < vis ? (int)((fb >> (6 * i)) & 63u) : -1;
  t.deck_pos = 2 * AP + (t.stage == 0 ? 0 : t.stage == 1 ? 4 : t.stage == 2 ? 6 : 8);
}

template <int MAXP>
PULSE_HD void packed_store(const PokerTable<MAXP>& t, const PokerTable<MAXP>& o, const PulsePokerPacked& S,
                           long long g, long long n, int AP) {
  uint32_t m1 = pkp_meta(t.stage, t.idx, t.agg, t.button, t.acted, t.done, t.eq_stage);
  uint32_t m0 = pkp_meta(o.stage, o.idx, o.agg, o.button, o.acted, o.done, o.eq_stage);
  if (m1 != m0 || t.st != o.st || t.pot != o.pot || t.highest != o.highest)
    pkp_st4(S.tbl + 4 * g, m1, t.st, (uint32_t)t.pot, (uint32_t)t.highest);
  if (t.last_raise != o.last_raise) S.tbl2[2 * g] = (uint32_t)t.last_raise;
#pragma unroll
  for (int s = 0; s < MAXP; ++s)
    if (s < AP) {
      long long off = (long long)s * n + g;
      if (t.stack[s] != o.stack[s]) S.stack[off] = t.stack[s];
      if (t.crb[s] != o.crb[s]) S.crb[off] = t.crb[s];
      if (t.inv[s] != o.inv[s]) S.inv[off] = t.inv[s];
    }
}

// Scratch for the partial Fisher-Yates shuffle: 52 card bytes addressed by a run-time index.  In a
// thread's local memory such an array costs one L1 sector per lane per access (lanes index different
// bytes) -- the reset kernel spent 28% of its stall samples on STL/LDL.  PackedDeckShared keeps it in
// shared memory, word-interleaved across the block's threads (byte i of thread t lives in word
// (i / 4) * block + t), so any index pattern across a warp is bank-conflict free.
struct PackedDeckLocal {
  uint8_t d[52];
  PULSE_HD int get(int i) const { return d[i]; }
  PULSE_HD void set(int i, int v) { d[i] = (uint8_t)v; }
  PULSE_HD void fill_identity() { for (int i = 0; i < 52; ++i) d[i] = (uint8_t)(i + 1); }
};
struct PackedDeckShared {
  uint8_t* base;     // this thread's byte 0 of word 0
  int word_stride;   // bytes between consecutive words of one thread: 4 * blockDim.x
  PULSE_HD int get(int i) const { return base[(i >> 2) * word_stride + (i & 3)]; }
  PULSE_HD void set(int i, int v) { base[(i >> 2) * word_stride + (i & 3)] = (uint8_t)v; }
  PULSE_HD void fill_identity() {
#pragma unroll
    for (int w = 0; w < 13; ++w)  // cards 4w+1 .. 4w+4, little endian
      *reinterpret_cast<uint32_t*>(base + w * word_stride) = 0x04030201u + 0x04040404u * (uint32_t)w;
  }
};

// Deals a new hand to one table (PokerGPU.py:73-157 for a single row).  Cards come from
// `prefixed` (int32 [N,52]) or from the chance tape; only the first 2AP+8 deck positions matter.
// keep_stacks == 0 fills every stack with starting_bbs; otherwise busted / oversized stacks are
// refilled and rolled by `rotation`.
template <int MAXP, class DeckBuf>
PULSE_HD void packed_deal(PokerTable<MAXP>& t, uint32_t& fb, const PulsePokerPacked& S, const PokerParams& pr,
                          const int32_t* __restrict__ prefixed, uint64_t seed, uint64_t env_id, uint32_t hand,
                          int rotation, int keep_stacks, int button, int starting_bbs, int max_bbs, long long g,
                          long long n, DeckBuf& dk) {
  const int P = pr.P, AP = pr.AP;
  const int need = 2 * AP + 8;
  if (prefixed) {
    for (int i = 0; i < need; ++i) dk.set(i, prefixed[g * 52 + i]);
  } else {
    dk.fill_identity();
    PulseU4 blk;
    blk.x = blk.y = blk.z = blk.w = 0;
    for (int i = 0; i < need; ++i) {  // forward Fisher-Yates fixes position i at iteration i
      if ((i & 3) == 0) blk = pulse_tape(seed, env_id, PULSE_STREAM_POKER_DECK, hand * 16u + (uint32_t)(i >> 2));
      uint32_t u = (i & 3) == 0 ? blk.x : (i & 3) == 1 ? blk.y : (i & 3) == 2 ? blk.z : blk.w;
      int j = i + (int)pulse_mulhi32(u, (uint32_t)(52 - i));
      const int ci = dk.get(i), cj = dk.get(j);
      dk.set(j, ci); dk.set(i, cj);
    }
  }
  // the cards this deal uses: two per dealt seat, then burn/flop/burn/turn/burn/river
  const int b0 = 2 * AP;
  const int f0 = dk.get(b0 + 1), f1 = dk.get(b0 + 2), f2 = dk.get(b0 + 3), tn = dk.get(b0 + 5), rv = dk.get(b0 + 7);
  int32_t old[MAXP];
#pragma unroll
  for (int s = 0; s < MAXP; ++s) {
    old[s] = starting_bbs;
    if (s < P && keep_stacks) {
      int v = S.stack[(long long)s * n + g];
      old[s] = (v == 0 || v > max_bbs) ? starting_bbs : v;  // :105-108
    }
  }
  uint32_t st = 0xFFFFFFFFu;
#pragma unroll
  for (int s = 0; s < MAXP; ++s) {
    int src = (s - rotation) % P;  // torch.roll (:109-110)
    if (src < 0) src += P;
    t.stack[s] = s < P ? pk_sel(old, keep_stacks ? src : s) : 0;
    t.crb[s] = 0; t.inv[s] = 0;
    const bool dealt = s < AP;
    t.hole0[s] = dealt ? dk.get(2 * s) : -1;
    t.hole1[s] = dealt ? dk.get(2 * s + 1) : -1;
    if (s < P) st = pk_set_status(st, s, dealt ? PK_ACTIVE : PK_SITOUT);
  }
  t.st = st;
  const int bb = AP == 2 ? (button + 1) % AP : (button + 2) % AP;  // :123-128
  int bstack = pk_sel(t.stack, bb) - 1;                            // post_blinds :188-199
  pk_put(t.stack, bb, bstack); pk_put(t.crb, bb, 1); pk_put(t.inv, bb, 1);
  if (bstack == 0) t.st = pk_set_status(t.st, bb, PK_ALLIN);
  t.pot = 1; t.highest = 1; t.last_raise = 1;
#pragma unroll
  for (int i = 0; i < 5; ++i) t.board[i] = -1;
  t.stage = 0; t.idx = AP == 2 ? button : (bb + 1) % AP;           // :130-133
  t.agg = bb; t.acted = 0; t.button = button; t.deck_pos = 2 * AP; t.eq_stage = 0; t.done = false; t.dirty = false;
  fb = (uint32_t)f0 | ((uint32_t)f1 << 6) | ((uint32_t)f2 << 12) | ((uint32_t)tn << 18) | ((uint32_t)rv << 24);
  // hand values of every dealt seat on the flop, turn and river boards of this deal
  const uint64_t flop = pe_add_card64(f2, pe_add_card64(f1, pe_add_card64(f0, 0)));
  const uint64_t turn = pe_add_card64(tn, flop), river = pe_add_card64(rv, turn);
  for (int s = 0; s < AP; ++s) {  // rolled: one copy of the evaluator, hole cards re-read from the scratch
    const uint64_t h = pe_add_card64(dk.get(2 * s + 1), pe_add_card64(dk.get(2 * s), 0));
    const int v5 = pe_eval_mask64(h | flop), v6 = pe_eval_mask64(h | turn), v7 = pe_eval_mask64(h | river);
    const int f = (v5 >= 0 && v5 < S.n_hand_ranks) ? PULSE_LDG(S.hand_ranks + v5) : 0;  // HR[HR[p5]] (:521)
    S.ev[(long long)s * n + g] = (uint64_t)(uint32_t)v7 | ((uint64_t)(uint32_t)v6 << 16) | ((uint64_t)(uint32_t)f << 32);
  }
}

// equity of a seat at `stage` from its stored hand-value word (PokerGPU.py:455-525)
PULSE_HD float pkp_equity_from(uint64_t ev, int stage) {
  if (stage < 1 || stage > 3) return 0.5f;  // :461
  const int v = stage == 3 ? (int)(ev & 0xFFFFu) : stage == 2 ? (int)((ev >> 16) & 0xFFFFu) : (int)(ev >> 32);
  const float lo = stage == 1 ? 74359.0f : 4109.0f, den = stage == 1 ? 749420.0f : 32765.0f;  // :15-18, :65-66
  const float e = PULSE_FDIV(PULSE_FSUB((float)v, lo), den);
  return fminf(fmaxf(e, 0.0f), 1.0f);
}

// Hand-value words of one table.  PRE == true (heads-up): both seats' words are requested together
// with the rest of the table, so neither the actor's equity nor the showdown waits for a second,
// dependent trip to DRAM (+8 B read per table-step).  PRE == false: one word is fetched when it is
// needed (the actor's, or each live seat's at showdown) -- with six or ten seats prefetching all of
// them would cost more bandwidth than the latency is worth.
template <int MAXP, bool PRE>
struct PackedEv;
template <int MAXP>
struct PackedEv<MAXP, true> {
  const uint64_t* ev; long long g, n;
  uint64_t w[MAXP];
  PULSE_HD void prefetch(int AP) {
#ifdef __CUDA_ARCH__
    // part of the step's ordered first-level loads (see packed_load): no branch, seats beyond AP re-read seat 0
#pragma unroll
    for (int s = 0; s < MAXP; ++s) {
      const uint64_t v = *reinterpret_cast<const volatile uint64_t*>(ev + (long long)(s < AP ? s : 0) * n + g);
      w[s] = s < AP ? v : 0;
    }
#else
    for (int s = 0; s < MAXP; ++s) w[s] = s < AP ? ev[(long long)s * n + g] : 0;
#endif
  }
  PULSE_HD uint64_t word(int s) const {
    uint64_t v = w[0];
#pragma unroll
    for (int k = 1; k < MAXP; ++k) if (s == k) v = w[k];
    return v;
  }
  template <int MP>
  PULSE_HD int32_t rank(const PokerTable<MP>&, int s) const { return (int32_t)(word(s) & 0xFFFFu); }  // showdown rank = seven-card value
};
template <int MAXP>
struct PackedEv<MAXP, false> {
  const uint64_t* ev; long long g, n;
  PULSE_HD void prefetch(int) {}
  PULSE_HD uint64_t word(int s) const { return ev[(long long)s * n + g]; }
  template <int MP>
  PULSE_HD int32_t rank(const PokerTable<MP>&, int s) const { return (int32_t)(word(s) & 0xFFFFu); }
};

// unconditional store of a freshly dealt table
template <int MAXP>
PULSE_HD void packed_store_all(const PokerTable<MAXP>& t, uint32_t fb, const PulsePokerPacked& S, long long g,
                               long long n, int P) {
  pkp_st4(S.tbl + 4 * g, pkp_meta(t.stage, t.idx, t.agg, t.button, t.acted, t.done, t.eq_stage), t.st, (uint32_t)t.pot,
          (uint32_t)t.highest);
  pkp_st2(S.tbl2 + 2 * g, (uint32_t)t.last_raise, fb);
  if (S.idx) S.idx[g] = t.idx;
  if (S.prev_invested) S.prev_invested[g] = 0;   // :151
#pragma unroll
  for (int s = 0; s < MAXP; ++s)
    if (s < P) {
      long long off = (long long)s * n + g;
      S.stack[off] = t.stack[s]; S.crb[off] = t.crb[s]; S.inv[off] = t.inv[s];
      S.hole[off] = (uint16_t)((t.hole0[s] & 63) | ((t.hole1[s] & 63) << 6));
    }
}

// Scripted seats evaluated inside the step (rules in poker_policy.cuh): the acting seat's
// kind picks the rule, its hole cards and the pot are already in registers.
struct PackedPolicy {
  PulseSeatKinds kinds;
  const uint32_t* chance2;   // uint32 [N,2] injected draws, or NULL: tape (seed, env, PULSE_STREAM_ACTION, counter)
  int64_t* chosen;           // int64 [N] action played, or NULL
  int32_t* seat_before;      // int32 [N] seat that acted (info["seat_idx"] before the step), or NULL
  uint32_t counter;
};

// One step of one table in the packed layout (PokerGPU.py:527-633).  auto_reset: a table that was
// already done when the step began is dealt a new hand instead (engine extension):
// the action is ignored, reward 0, done 0 and the observation is the new hand's first one.
// AUTO: what a table that was already done when the step began does -- 0 nothing (stays done), 1 deal inline
// (host emulation of the semantics), 2 mark for the redeal pass that follows on the same stream (device).  A compile-time choice so the plain
// step kernel carries no shuffle / evaluator code at all.
// Result of packed_step_core: what the step did to one table, for the caller to write out (float32 observation row
// + reward + done arrays, or one compact wire row).
struct PackedStepResult {
  float reward;
  bool was_done;   // the table was already finished when the step began (PokerGPU.py:530)
  bool done;       // done flag to report
};

// The step proper on a table held in registers (PokerGPU.py:531-632): no loads, no stores.  evw hands out the seats'
// per-deal hand-value words; prev_inv = the actor's current_round_bet before the action (:539).
template <int MAXP, class Ev>
PULSE_HD PackedStepResult packed_step_regs(PokerTable<MAXP>& t, uint32_t fb, const Ev& evw, const PokerParams& pr,
                                           long long a, int& prev_inv) {
  const PackedDeck deck{fb, 2 * pr.AP};
  const bool prev_done = t.done;
  const int actor = t.idx;                                           // :531
  const bool legal = pk_status(t.st, actor) == PK_ACTIVE && !prev_done;  // :532-537
  prev_inv = pk_sel(t.crb, actor);                                   // :539
  // the reference refreshes its equity cache here when the street has changed since the last step (:542);
  // the packed layout only remembers for which street the cache stands
  if (t.stage >= 1 && t.stage <= 3) t.eq_stage = t.stage;
  float e = 0.5f;
  if (legal)  // the reward is zeroed otherwise (:632), so the equity is only needed for legal actors
    e = (t.stage >= 1 && t.stage <= 3) ? pkp_equity_from(evw.word(actor), t.stage) : 0.5f;
  pk_execute(t, a);                                                  // :546
  pk_advance(t, pr, deck);                                           // :547-616
  if (t.done && !prev_done) {                                        // :619-622
    pk_fold_winner(t, pr);
    pk_showdown(t, pr, deck, evw);
  }
  if (t.done) {                                                      // :625-628
#pragma unroll
    for (int s = 0; s < MAXP; ++s) { t.crb[s] = 0; t.inv[s] = 0; }
    t.highest = 0;
  }
  const float reward = legal ? pk_reward(t, pr, a, e, prev_inv) : 0.0f;   // :631-632
  return PackedStepResult{reward, prev_done, t.done};
}

template <int MAXP, bool POLICY = false, int AUTO = 1, class ActT = int64_t>
PULSE_HD PackedStepResult packed_step_core(const PulsePokerPacked& S, const PokerParams& pr, const ActT* __restrict__ actions,
                                           int auto_reset, uint64_t seed, uint64_t env_base, int starting_bbs, int max_bbs,
                                           long long g, long long n, const PackedPolicy* pol, PokerTable<MAXP>& t) {
  const int P = pr.P, AP = pr.AP;
  uint32_t fb;
  PackedEv<MAXP, (MAXP <= 2)> evw{S.ev, g, n};
  evw.prefetch(AP);
  long long a = 0;
#ifdef __CUDA_ARCH__
  if (!POLICY) a = (long long)*reinterpret_cast<const volatile ActT*>(actions + g);  // first of the ordered first-level loads
#else
  if (!POLICY) a = (long long)actions[g];
#endif
  if (sizeof(ActT) == 1 && a == 255) a = -1;   // byte-wide action ids: 255 stands for "negative" (matches no branch)
  packed_load(t, fb, S, g, n, AP);
  if (POLICY) {  // build_actions: computed from the pre-step state for every table, done or not (utils.py:108-123)
    const int kind = pol->kinds.kind[t.idx & 15];
    if (kind == PULSE_AGENT_EXTERNAL) {
      a = (long long)actions[g];
    } else {
      uint32_t w0, w1;
      if (pol->chance2) { w0 = pol->chance2[2 * g]; w1 = pol->chance2[2 * g + 1]; }
      else { PulseU4 b = pulse_tape(seed, env_base + (uint64_t)g, PULSE_STREAM_ACTION, pol->counter); w0 = b.x; w1 = b.y; }
      a = pulse_policy_action(kind, pk_sel(t.hole0, t.idx), pk_sel(t.hole1, t.idx), (float)t.pot, w0, w1);
    }
    if (pol->chosen) pol->chosen[g] = a;
    if (pol->seat_before) pol->seat_before[g] = t.idx;
  }
  if (AUTO == 2 && auto_reset && t.done) {
    // device path: the deal is left to the redeal pass that follows the step kernel on the same stream
    // (poker_packed.cu).  Dealing here would put the shuffle and 3 x AP hand evaluations into the step
    // kernel's register budget and make every warp with one finished table wait for it (measured: 1.7 ms per
    // 4M-table step when half the tables redeal, against 0.16 + 0.2 ms for step + redeal pass).
    S.hand_ctr[g] |= PKP_REDEAL_MARK;
    return PackedStepResult{0.0f, true, false};   // the row of the finished hand is a placeholder the redeal pass overwrites
  }
  if (AUTO == 1 && auto_reset && t.done) {
    uint32_t hand = S.hand_ctr[g] + 1u;
    S.hand_ctr[g] = hand;
    PackedDeckLocal dk;
    packed_deal(t, fb, S, pr, nullptr, seed, env_base + (uint64_t)g, hand, 0, 1, t.button + 1 < AP ? t.button + 1 : (t.button + 1) % AP,
                starting_bbs, max_bbs, g, n, dk);
    packed_store_all(t, fb, S, g, n, P);
    return PackedStepResult{0.0f, true, false};
  }
  const PokerTable<MAXP> t0 = t;
  int prev_inv = 0;
  const PackedStepResult r = packed_step_regs<MAXP>(t, fb, evw, pr, a, prev_inv);
  if (S.prev_invested) S.prev_invested[g] = prev_inv;
  packed_store(t, t0, S, g, n, AP);
  if (S.idx && t.idx != t0.idx) S.idx[g] = t.idx;
  return r;
}

// The step with the reference's own outputs: float32 observation row (:633), reward and done arrays.
template <int MAXP, bool POLICY = false, int AUTO = 1>
PULSE_HD void packed_step_one(const PulsePokerPacked& S, const PokerParams& pr, const int64_t* __restrict__ actions,
                              float* __restrict__ obs_row, float* __restrict__ reward_out,
                              uint8_t* __restrict__ done_out, int auto_reset, uint64_t seed, uint64_t env_base,
                              int starting_bbs, int max_bbs, long long g, long long n,
                              const PackedPolicy* pol = nullptr) {
  PokerTable<MAXP> t;
  const PackedStepResult r = packed_step_core<MAXP, POLICY, AUTO>(S, pr, actions, auto_reset, seed, env_base,
                                                                  starting_bbs, max_bbs, g, n, pol, t);
  reward_out[g] = r.reward;
  done_out[g] = r.done;
  pk_obs(t, pr, obs_row);
}

template <int MAXP, class DeckBuf>
PULSE_HD void packed_reset_one(const PulsePokerPacked& S, const PokerParams& pr, const int32_t* __restrict__ prefixed,
                               uint64_t seed, uint64_t env_base, uint32_t hand, int rotation, int first_reset,
                               int starting_bbs, int max_bbs, float* __restrict__ obs_row, long long g, long long n,
                               DeckBuf& dk) {
  PokerTable<MAXP> t;
  uint32_t fb;
  int button = 0;
  if (!(first_reset & 2)) button = (int)(((S.tbl[4 * g] >> 11) & 15u) + 1u) % pr.AP;  // :121
  // the tape's hand index never goes back for a table: whole-batch resets mixed with per-table redeals (which
  // advance hand_ctr on their own) must not replay a deck.  first_reset bit 2 restarts the index at `hand`.
  if (S.hand_ctr && !(first_reset & 4)) {
    const uint32_t nxt = (S.hand_ctr[g] & ~PKP_REDEAL_MARK) + 1u;
    if (nxt > hand) hand = nxt;
  }
  packed_deal(t, fb, S, pr, prefixed, seed, env_base + (uint64_t)g, hand, rotation, !(first_reset & 1), button,
              starting_bbs, max_bbs, g, n, dk);
  packed_store_all(t, fb, S, g, n, pr.P);
  if (S.hand_ctr) S.hand_ctr[g] = hand;
  if (obs_row) pk_obs(t, pr, obs_row);
}

// The redeal pass of auto_reset: deals the next hand to a table the step kernel marked.  Returns whether it did.
template <int MAXP, class DeckBuf>
PULSE_HD bool packed_redeal_one(const PulsePokerPacked& S, const PokerParams& pr, uint64_t seed, uint64_t env_base,
                                int starting_bbs, int max_bbs, float* __restrict__ obs_row, long long g, long long n,
                                DeckBuf& dk) {
  const uint32_t ctr = S.hand_ctr[g];
  if (!(ctr & PKP_REDEAL_MARK)) return false;
  packed_reset_one<MAXP>(S, pr, nullptr, seed, env_base, (ctr & ~PKP_REDEAL_MARK) + 1u, 0, 4, starting_bbs, max_bbs,
                         obs_row, g, n, dk);   // stores hand_ctr = new counter: the mark is cleared
  return true;
}

template <int MAXP>
PULSE_HD void packed_reset_one(const PulsePokerPacked& S, const PokerParams& pr, const int32_t* __restrict__ prefixed,
                               uint64_t seed, uint64_t env_base, uint32_t hand, int rotation, int first_reset,
                               int starting_bbs, int max_bbs, float* __restrict__ obs_row, long long g, long long n) {
  PackedDeckLocal dk;
  packed_reset_one<MAXP>(S, pr, prefixed, seed, env_base, hand, rotation, first_reset, starting_bbs, max_bbs, obs_row,
                         g, n, dk);
}

// reference-shaped tensors -> packed: the inverse of packed_unpack_one (SURVEY 8(b) pulse_poker_pack), for state that was
// edited through the reference's attributes (white-box tests, hand-written positions) and should go back onto the
// throughput layout.  The packed layout holds less than the tensors can express, so the function says whether the
// table was representable: deck position as the street implies (2 AP + 0 / 4 / 6 / 8), no board card beyond the
// street's visible ones, cards 1..52 for dealt seats, seat / status / street numbers in range.  The visible board is
// taken from `board` (authoritative after an edit), the cards still to come from `decks`; the per-deal hand-value
// words are recomputed; the equity cache's street is the one whose values the `equities` tensor holds (when given),
// else the street itself (one before it when equity_dirty).  hand_ctr is left alone.
template <int MAXP>
PULSE_HD bool packed_pack_one(const PulsePokerPacked& S, const PulsePokerState& C, const PokerParams& pr, long long g,
                              long long n) {
  const int P = pr.P, AP = pr.AP;
  bool ok = true;
  PokerTable<MAXP> t;
  const int stage = C.stages[g];
  ok &= stage >= 0 && stage <= 5;
  t.stage = stage & 7;
  t.idx = C.idx[g]; t.agg = C.agg[g]; t.button = C.button[g]; t.acted = C.acted[g];
  ok &= t.idx >= 0 && t.idx < AP && t.agg >= 0 && t.agg < 16 && t.button >= 0 && t.button < AP && t.acted >= 0;
  t.idx &= 15; t.agg &= 15; t.button &= 15;
  if (t.acted < 0) t.acted = 0;
  t.done = C.is_done[g] != 0;
  t.pot = C.pots[g]; t.highest = C.highest[g]; t.last_raise = C.last_raise_size[g];
  uint32_t st = 0xFFFFFFFFu;
#pragma unroll
  for (int s = 0; s < MAXP; ++s) {
    t.stack[s] = 0; t.crb[s] = 0; t.inv[s] = 0; t.hole0[s] = -1; t.hole1[s] = -1;
    if (s < P) {
      const long long off = g * P + s;
      t.stack[s] = C.stacks[off]; t.crb[s] = C.current_round_bet[off]; t.inv[s] = C.total_invested[off];
      const int v = C.status[off];
      ok &= v >= 0 && v <= 3 && (s < AP || v == PK_SITOUT);
      st = pk_set_status(st, s, v & 3);
      const int h0 = C.hands[2 * off], h1 = C.hands[2 * off + 1];
      if (s < AP) { ok &= h0 >= 1 && h0 <= 52 && h1 >= 1 && h1 <= 52; t.hole0[s] = h0; t.hole1[s] = h1; }
      else ok &= h0 == -1 && h1 == -1;
    }
  }
  t.st = st;
  const int vis = stage == 0 ? 0 : stage == 1 ? 3 : stage == 2 ? 4 : 5;
  const int dpos = 2 * AP + (stage == 0 ? 0 : stage == 1 ? 4 : stage == 2 ? 6 : 8);
  ok &= C.deck_positions[g] == dpos;
  const int32_t* deck = C.decks + g * 52;
  const int b0 = 2 * AP;
  int card[5] = {deck[b0 + 1], deck[b0 + 2], deck[b0 + 3], deck[b0 + 5], deck[b0 + 7]};
#pragma unroll
  for (int i = 0; i < 5; ++i) {
    const int b = C.board[g * 5 + i];
    if (i < vis) { ok &= b >= 1 && b <= 52; card[i] = b; }
    else ok &= b == -1;
    ok &= card[i] >= 1 && card[i] <= 52;
    card[i] &= 63;
  }
  const uint32_t fb = (uint32_t)card[0] | ((uint32_t)card[1] << 6) | ((uint32_t)card[2] << 12) | ((uint32_t)card[3] << 18) |
                      ((uint32_t)card[4] << 24);
  // hand values of every dealt seat on the flop, turn and river boards (as packed_deal)
  const uint64_t flop = pe_add_card64(card[2], pe_add_card64(card[1], pe_add_card64(card[0], 0)));
  const uint64_t turn = pe_add_card64(card[3], flop), river = pe_add_card64(card[4], turn);
  for (int s = 0; s < AP; ++s) {
    const uint64_t h = pe_add_card64(pk_sel(t.hole1, s), pe_add_card64(pk_sel(t.hole0, s), 0));
    const int v5 = pe_eval_mask64(h | flop), v6 = pe_eval_mask64(h | turn), v7 = pe_eval_mask64(h | river);
    const int f = (v5 >= 0 && v5 < S.n_hand_ranks) ? PULSE_LDG(S.hand_ranks + v5) : 0;
    S.ev[(long long)s * n + g] = (uint64_t)(uint32_t)v7 | ((uint64_t)(uint32_t)v6 << 16) | ((uint64_t)(uint32_t)f << 32);
  }
  // the street the reference's equity cache stands for
  int eq = 0;
  if (stage >= 1 && stage <= 3) eq = (C.equity_dirty && C.equity_dirty[g]) ? stage - 1 : stage;
  else if (stage >= 4) eq = 3;
  if (C.equities) {
    int found = -1;
    for (int k = 0; k < 5 && found < 0; ++k) {
      const int cand = k == 0 ? eq : k - 1;
      bool same = true;
      for (int s = 0; s < AP; ++s) {
        const float want = cand == 0 ? 0.5f : pkp_equity_from(S.ev[(long long)s * n + g], cand);
        same &= C.equities[g * AP + s] == want;
      }
      if (same) found = cand;
    }
    if (found >= 0) eq = found;
  }
  t.eq_stage = eq;
  t.deck_pos = dpos; t.dirty = false;
#pragma unroll
  for (int i = 0; i < 5; ++i) t.board[i] = i < vis ? card[i] : -1;
  packed_store_all(t, fb, S, g, n, P);
  if (S.prev_invested) S.prev_invested[g] = C.prev_invested ? C.prev_invested[g] : 0;
  return ok;
}

// ---- enumerated hand strength ---------------------------------------------------------
// The reference's equity is a table-walk proxy that its own notes call a known weakness
// (/root/reference/rl/depth_notes.txt:302-333): on the turn and flop it feeds 6- and 5-card table values through
// constants fitted to other quantities (PokerGPU.py:497-523).  This is the enumerated form of the same river
// quantity: the expectation, over the board cards still to come, of the river strength
// clamp((v7 - 4109) / 32765, 0, 1) (:477-481), from the seat's own point of view (unseen = 52 minus its two hole
// cards and the visible board).  River: the reference's value itself.  Turn: 46 river cards.  Flop: C(47,2) = 1081
// turn/river pairs.  Preflop: 0.5, as the reference (:461).  The numerators are summed as integers, so the result
// does not depend on the enumeration order: strength = float(sum) / (float(count) * 32765).
PULSE_HD int64_t pkp_strength_num(uint64_t m7) {
  int v = pe_eval_mask64(m7) - 4109;
  return v < 0 ? 0 : (v > 32765 ? 32765 : v);
}
PULSE_HD float packed_enumerated_strength(int stage, int h0, int h1, uint32_t fb) {
  if (stage < 1 || stage > 3) return 0.5f;
  uint64_t known = pe_add_card64(h1, pe_add_card64(h0, 0));
  const int vis = stage == 1 ? 3 : stage == 2 ? 4 : 5;
  for (int i = 0; i < vis; ++i) known = pe_add_card64((int)((fb >> (6 * i)) & 63u), known);
  int64_t sum = 0, count = 0;
  if (stage == 3) { sum = pkp_strength_num(known); count = 1; }
  else {
    for (int c1 = 1; c1 <= 52; ++c1) {
      const uint64_t m1 = pe_add_card64(c1, 0);
      if (m1 & known) continue;
      if (stage == 2) { sum += pkp_strength_num(known | m1); ++count; continue; }
      for (int c2 = c1 + 1; c2 <= 52; ++c2) {
        const uint64_t m2 = pe_add_card64(c2, 0);
        if (m2 & known) continue;
        sum += pkp_strength_num(known | m1 | m2); ++count;
      }
    }
  }
  return PULSE_FDIV((float)sum, PULSE_FMUL((float)count, 32765.0f));
}

// ---- showdown equity (second half) -------------------------------------------------------
// What the reference's notes ask for and call out of reach: equity "using ALL players card info (not just 100%
// accuracy with current players hand)" (/root/reference/rl/depth_notes.txt:316-321).  The env knows every seat's hole
// cards, so this is the all-in equity of the table as it stands: over every completion of the board from the cards
// that are in nobody's hand (folded seats' cards are dead) and not on the visible board, the share of the pot a
// contender (status ACTIVE or ALLIN) would win at showdown -- 1 for a sole best seven-card value, 1/k in a k-way
// tie, by the same evaluator values the settlement compares (poker_core.cuh::pk_showdown).
//   river / settled hand: the one board;  turn: every river card (52 - 2 AP - 4 boards);  flop: every turn/river
//   pair (heads-up C(45,2) = 990, ten seats C(29,2) = 406);  preflop: C(52 - 2 AP, 5) is 1.7 M boards heads-up, so
//   `samples` boards are drawn instead, board q from one block of the chance tape at (seed, table id, counter,
//   stream PULSE_STREAM_EQUITY | q << 4): a forward Fisher-Yates over the unseen cards in (suit, rank) order.
// Shares are summed as integers in units of 1/2520 (lcm of 1..10), so the result does not depend on how the boards
// are split over lanes: equity = float(double(sum) / (double(boards) * 2520)).  Folded / sitting-out seats: 0.
constexpr int PKQ_UNIT = 2520;
constexpr int PKQ_MAX_SAMPLES = 65536;

template <int MAXP>
PULSE_HD void pkq_score(const uint64_t (&hole)[MAXP], uint32_t contenders, uint64_t board, int32_t (&num)[MAXP]) {
  int best = -1;
  uint32_t win = 0;
#pragma unroll
  for (int s = 0; s < MAXP; ++s) {
    if ((contenders >> s) & 1u) {
      const int v = pe_eval_mask64(hole[s] | board);
      if (v > best) { best = v; win = 1u << s; }
      else if (v == best) win |= 1u << s;
    }
  }
  const int k = PULSE_POPC(win);
  const int share = PKQ_UNIT / (k > 0 ? k : 1);
#pragma unroll
  for (int s = 0; s < MAXP; ++s) num[s] += ((win >> s) & 1u) ? share : 0;
}

// Adds to num[] the shares of the boards with index == lane (mod nl); returns the number of boards of ALL lanes.
template <int MAXP>
PULSE_HD int packed_showdown_equity_part(const PokerTable<MAXP>& t, int AP, int samples, uint64_t seed, uint64_t env_id,
                                         uint32_t counter, int lane, int nl, int32_t (&num)[MAXP]) {
  uint64_t hole[MAXP];
  uint64_t dead = 0;
  uint32_t contenders = 0;
#pragma unroll
  for (int s = 0; s < MAXP; ++s) {
    hole[s] = 0;
    if (s < AP) {
      hole[s] = pe_add_card64(t.hole1[s], pe_add_card64(t.hole0[s], 0));
      dead |= hole[s];
      const int st = pk_status(t.st, s);
      if (st == PK_ACTIVE || st == PK_ALLIN) contenders |= 1u << s;
    }
  }
  uint64_t board = 0;
  int vis = 0;
#pragma unroll
  for (int i = 0; i < 5; ++i)
    if (t.board[i] >= 1) { board = pe_add_card64(t.board[i], board); ++vis; }
  dead |= board;
  if (PULSE_POPC(contenders) <= 1) {            // everybody else folded: every board is his
    if (lane == 0) {
#pragma unroll
      for (int s = 0; s < MAXP; ++s) num[s] += ((contenders >> s) & 1u) ? PKQ_UNIT : 0;
    }
    return 1;
  }
  if (vis == 5) {
    if (lane == 0) pkq_score<MAXP>(hole, contenders, board, num);
    return 1;
  }
  uint8_t pos[52];                              // bit positions (16 suit + rank) of the unseen cards, ascending
  int m = 0;
  for (int b = 0; b < 64; ++b)
    if ((b & 15) < 13 && !((dead >> b) & 1ull)) pos[m++] = (uint8_t)b;
  if (vis == 4) {
    for (int i = lane; i < m; i += nl) pkq_score<MAXP>(hole, contenders, board | (1ull << pos[i]), num);
    return m;
  }
  if (vis == 3) {
    // pair (i, j), i < j, in row-major order; row i holds m - 1 - i pairs; this lane owns every nl-th pair
    int i = 0, r = lane;
    while (i < m - 1 && r >= m - 1 - i) { r -= m - 1 - i; ++i; }
    while (i < m - 1) {
      pkq_score<MAXP>(hole, contenders, board | (1ull << pos[i]) | (1ull << pos[i + 1 + r]), num);
      r += nl;
      while (i < m - 1 && r >= m - 1 - i) { r -= m - 1 - i; ++i; }
    }
    return m * (m - 1) / 2;
  }
  // preflop (vis == 0): sampled boards
  for (int q = lane; q < samples; q += nl) {
    const PulseU4 w = pulse_tape(seed, env_id, PULSE_STREAM_EQUITY | ((uint32_t)q << 4), counter);
    const uint32_t a = pulse_mulhi32(w.x, (uint32_t)(m * (m - 1)));
    int pick[5];
    pick[0] = (int)(a / (uint32_t)(m - 1));
    pick[1] = 1 + (int)(a % (uint32_t)(m - 1));
    pick[2] = 2 + (int)pulse_mulhi32(w.y, (uint32_t)(m - 2));
    pick[3] = 3 + (int)pulse_mulhi32(w.z, (uint32_t)(m - 3));
    pick[4] = 4 + (int)pulse_mulhi32(w.w, (uint32_t)(m - 4));
    uint64_t bd = 0;
    for (int k = 0; k < 5; ++k) {               // forward Fisher-Yates: slot k takes the card at slot pick[k] >= k
      const uint8_t x = pos[pick[k]]; pos[pick[k]] = pos[k]; pos[k] = x;
      bd |= 1ull << x;
    }
    pkq_score<MAXP>(hole, contenders, bd, num);
    for (int k = 4; k >= 0; --k) {              // undo, so the next sample starts from the sorted list again
      const uint8_t x = pos[k]; pos[k] = pos[pick[k]]; pos[pick[k]] = x;
    }
  }
  return samples;
}

PULSE_HD float pkq_equity(int64_t sum, int boards) {
  return (float)((double)sum / ((double)boards * (double)PKQ_UNIT));
}

// packed -> reference-shaped tensors: the mirror behind attribute access on the drop-in env class, and the
// inspection path of the parity tests.  Every pointer of C may be NULL (field not wanted).  decks: re-drawn from the
// tape (or copied from the prefixed decks of the last reset) when `with_decks`; equities / equity_dirty: the
// reference's cache as it stands (header of include/pulse_envs.h); prev_invested: the engine's own mirror, else 0.
template <int MAXP>
PULSE_HD void packed_unpack_one(const PulsePokerPacked& S, const PulsePokerState& C, const PokerParams& pr,
                                long long g, long long n, bool with_decks = false, const int32_t* prefixed = nullptr,
                                uint64_t seed = 0, uint64_t env_base = 0, bool fresh = false) {
  const int P = pr.P, AP = pr.AP;
  PokerTable<MAXP> t;
  uint32_t fb;
  packed_load(t, fb, S, g, n, AP);
  if (t.stage == 5) {
#pragma unroll
    for (int i = 0; i < 5; ++i) t.board[i] = (int)((fb >> (6 * i)) & 63u);
  }
  for (int s = 0; s < P; ++s) {
    long long off = g * P + s;
    if (C.stacks) C.stacks[off] = s < AP ? pk_sel(t.stack, s) : S.stack[(long long)s * n + g];
    if (C.current_round_bet) C.current_round_bet[off] = pk_sel(t.crb, s);
    if (C.total_invested) C.total_invested[off] = pk_sel(t.inv, s);
    if (C.status) C.status[off] = s < P ? pk_status(t.st, s) : PK_SITOUT;
    if (C.hands) { C.hands[2 * off] = pk_sel(t.hole0, s); C.hands[2 * off + 1] = pk_sel(t.hole1, s); }
  }
  if (C.board) for (int i = 0; i < 5; ++i) C.board[g * 5 + i] = t.board[i];
  if (C.pots) C.pots[g] = t.pot;
  if (C.stages) C.stages[g] = t.stage;
  if (C.highest) C.highest[g] = t.highest;
  if (C.last_raise_size) C.last_raise_size[g] = t.last_raise;
  if (C.idx) C.idx[g] = t.idx;
  if (C.agg) C.agg[g] = t.agg;
  if (C.acted) C.acted[g] = t.acted;
  if (C.button) C.button[g] = t.button;
  if (C.deck_positions) C.deck_positions[g] = t.deck_pos;
  if (C.is_done) C.is_done[g] = t.done;
  {  // blinds as reset placed them (:123-128)
    const int b = t.button;
    if (C.sb) C.sb[g] = AP == 2 ? b : (b + 1) % AP;
    if (C.bb) C.bb[g] = AP == 2 ? (b + 1) % AP : (b + 2) % AP;
  }
  if (C.equities)
    for (int s = 0; s < AP; ++s)
      C.equities[g * AP + s] = t.eq_stage == 0 ? 0.5f : pkp_equity_from(S.ev[(long long)s * n + g], t.eq_stage);
  if (C.equity_dirty) C.equity_dirty[g] = fresh || (t.stage >= 1 && t.stage <= 3 && t.stage != t.eq_stage);
  if (C.prev_invested) C.prev_invested[g] = S.prev_invested ? S.prev_invested[g] : 0;
  if (with_decks && C.decks) {
    int32_t* deck = C.decks + g * 52;
    if (prefixed) {
      for (int i = 0; i < 52; ++i) deck[i] = prefixed[g * 52 + i];
    } else {
      uint8_t d[52];
      pk_tape_deck(d, seed, env_base + (uint64_t)g, S.hand_ctr ? (S.hand_ctr[g] & ~PKP_REDEAL_MARK) : 0u);
      for (int i = 0; i < 52; ++i) deck[i] = d[i];
    }
  }
}
