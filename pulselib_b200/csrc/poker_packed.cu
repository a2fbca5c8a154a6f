// Packed-layout PokerGPU kernels and their C ABI (include/pulse_envs.h).  Layout and rationale:
// poker_packed_body.cuh.  HBM-bound: see DESIGN.md for bytes per table-step.
#include "poker_packed_body.cuh"
#include <stdlib.h>

namespace {

constexpr int kBlock = 128;

// Observation write-out.  Each thread builds its row in registers; the block stages the rows in
// shared memory and then writes its contiguous [kBlock, W] slab of obs with fully coalesced 16 B
// stores (a thread writing its own 64-160 B row straight to global would touch 32 different lines
// per store instruction).  MT > 0: max_players is the compile-time constant MT, W % 4 == 0, rows
// are staged as float4 with an odd row stride (in 16 B units) so both sides are conflict-free.
// MT == 0: any max_players, scalar staging with row stride W + 1.
// Ends with a block barrier so the staging buffer can be reused by the next grid-stride iteration.
template <int MT, int WT>
__device__ __forceinline__ void stage_obs_rows(const float (&row)[WT], float4* smem4, float* __restrict__ obs_out,
                                               long long base, long long g, long long n, int W) {
  const long long rows = (n - base) < kBlock ? (n - base) : kBlock;
  if (MT > 0) {
    constexpr int W4 = WT / 4, S4 = W4 | 1;
    if (g < n) {
#pragma unroll
      for (int c = 0; c < W4; ++c)
        smem4[threadIdx.x * S4 + c] = make_float4(row[4 * c], row[4 * c + 1], row[4 * c + 2], row[4 * c + 3]);
    }
    __syncthreads();
    float4* dst = reinterpret_cast<float4*>(obs_out + base * WT);
    const int total = (int)rows * W4;
#pragma unroll
    for (int it = 0; it < W4; ++it) {
      const int q = threadIdx.x + it * kBlock;
      if (q < total) {
        const int r = q / W4, c = q - r * W4;
        dst[q] = smem4[r * S4 + c];
      }
    }
  } else {
    float* smem = reinterpret_cast<float*>(smem4);
    float* mine = smem + threadIdx.x * (W + 1);
    if (g < n)
      for (int i = 0; i < W; ++i) mine[i] = row[i];
    __syncthreads();
    const int total = (int)rows * W;
    float* dst = obs_out + base * W;
    const int dq = kBlock / W, dr = kBlock - dq * W;
    int r = threadIdx.x / W, c = threadIdx.x - r * W;
    for (int e = threadIdx.x; e < total; e += kBlock) {
      dst[e] = smem[r * (W + 1) + c];
      r += dq; c += dr;
      if (c >= W) { c -= W; ++r; }
    }
  }
  __syncthreads();
}

// L2 prefetch one wave ahead.  The step is one short-lived CTA per 128-table tile and every first-level load
// of a tile goes to DRAM (ncu: long-scoreboard is the top stall, DRAM at 56% with six seats), so the kernel is
// bound by how many loads a resident thread set can keep in flight.  Each CTA therefore asks L2 (prefetch.global.L2:
// no register, no scoreboard) for the state lines of the tile `ahead` CTAs further on -- about one resident
// wave, i.e. the tile that takes this CTA's slot when it retires -- so that tile's loads meet L2 latency instead
// of DRAM latency.  DRAM traffic is unchanged (every line is still fetched once).
// Issued by the CTA's first warp only (the other three skip it with one uniform branch; spreading the lines over all
// 128 threads cost 85 instructions per warp in divergent address selection, 6.6% of an issue-bound kernel):
// one prefetch covers tbl (lanes 0-15), tbl2 (16-23) and the actions (24-31); then one prefetch per seat covers
// that seat's stack / crb / inv (4 lines each), hole (2) and, when asked, hand-value words (8) -- the lane's array,
// element size and line are chosen once, outside the seat loop.
template <int MAXP>
__device__ __forceinline__ void packed_prefetch_tile(const PulsePokerPacked& S, const int64_t* actions, int AP,
                                                     long long base, long long n, bool with_ev) {
  if (threadIdx.x >= 32 || base >= n) return;
  const int lane = threadIdx.x;
  const long long cnt = (n - base) < kBlock ? (n - base) : kBlock;
  {
    const char* p; int esz, k;
    if (lane < 16) { p = reinterpret_cast<const char*>(S.tbl + 4 * base); esz = 16; k = lane; }
    else if (lane < 24) { p = reinterpret_cast<const char*>(S.tbl2 + 2 * base); esz = 8; k = lane - 16; }
    else { p = reinterpret_cast<const char*>(actions + base); esz = actions ? 8 : 0; k = lane - 24; }
    if ((long long)k * 128 < cnt * esz) asm volatile("prefetch.global.L2 [%0];" ::"l"(p + k * 128));
  }
  const char* arr; int esz, k;
  if (lane < 4) { arr = reinterpret_cast<const char*>(S.stack); esz = 4; k = lane; }
  else if (lane < 8) { arr = reinterpret_cast<const char*>(S.crb); esz = 4; k = lane - 4; }
  else if (lane < 12) { arr = reinterpret_cast<const char*>(S.inv); esz = 4; k = lane - 8; }
  else if (lane < 14) { arr = reinterpret_cast<const char*>(S.hole); esz = 2; k = lane - 12; }
  else { arr = reinterpret_cast<const char*>(S.ev); esz = 8; k = lane - 14; }
  if (lane < (with_ev && S.ev ? 22 : 14) && (long long)k * 128 < cnt * esz) {
    arr += base * esz + k * 128;
    const long long seat_stride = n * esz;
#pragma unroll
    for (int s = 0; s < MAXP; ++s)
      if (s < AP) asm volatile("prefetch.global.L2 [%0];" ::"l"(arr + s * seat_stride));
  }
}

// Heads-up rows are 16 floats = 64 B = two 32-byte sectors: each thread writes its own row with two 256-bit stores
// (sm_100: st.global.v8.f32 -> STG.E.256).  Every store fills a whole sector, so nothing is gained by staging the
// rows through shared memory first -- and the staging (4 STS.128, two block barriers, 4 LDS.128, index arithmetic) was
// 18% of the instructions of an issue-bound kernel (profiles/r2b_autoreset_step_raw.csv).  dst must be 32-byte aligned.
__device__ __forceinline__ void store_row16(float* dst, const float (&r)[16]) {
  asm volatile("st.global.v8.f32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(dst), "f"(r[0]), "f"(r[1]), "f"(r[2]),
               "f"(r[3]), "f"(r[4]), "f"(r[5]), "f"(r[6]), "f"(r[7]) : "memory");
  asm volatile("st.global.v8.f32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(dst + 8), "f"(r[8]), "f"(r[9]), "f"(r[10]),
               "f"(r[11]), "f"(r[12]), "f"(r[13]), "f"(r[14]), "f"(r[15]) : "memory");
}

template <int MAXP, int MT, int MINB, bool POLICY = false, int AUTO = 0, bool DIRECT = false>
__global__ void __launch_bounds__(kBlock, MINB)
poker_packed_step_kernel(PulsePokerPacked S, PokerParams pr, const int64_t* __restrict__ actions,
                         float* __restrict__ obs_out, float* __restrict__ reward_out,
                         uint8_t* __restrict__ done_out, int auto_reset, uint64_t seed, uint64_t env_base,
                         int starting_bbs, int max_bbs, long long n, int l2_ahead,
                         PackedPolicy pol = PackedPolicy()) {
  extern __shared__ float4 smem4[];
  if (l2_ahead > 0)  // bit 30 of l2_ahead: include the hand-value words
    packed_prefetch_tile<MAXP>(S, actions, pr.AP, ((long long)blockIdx.x + (l2_ahead & 0xFFFFF)) * kBlock, n,
                               (l2_ahead >> 30) & 1);
  if (MT > 0) pr.M = MT;
  constexpr int WT = MT > 0 ? 13 + 3 * (MT - 1) : 13 + 3 * 15;  // row buffer size
  const int W = 13 + 3 * (pr.M - 1);
  for (long long base = (long long)blockIdx.x * kBlock; base < n; base += (long long)gridDim.x * kBlock) {
    const long long g = base + threadIdx.x;
    float row[WT];
    if (g < n)
      packed_step_one<MAXP, POLICY, AUTO>(S, pr, actions, row, reward_out, done_out, auto_reset, seed, env_base,
                                          starting_bbs, max_bbs, g, n, &pol);
    if (DIRECT && MT == 2) {
      if (g < n) store_row16(obs_out + g * 16, reinterpret_cast<const float(&)[16]>(row));
    } else {
      stage_obs_rows<MT, WT>(row, smem4, obs_out, base, g, n, W);
    }
  }
}

// Whole-batch deal.  Shared memory: the shuffle scratch (13 words per thread, word-interleaved, see
// PackedDeckShared) followed by the observation staging buffer of stage_obs_rows.
constexpr int kDeckSmemBytes = 13 * kBlock * 4;

template <int MAXP, int MT>
__global__ void __launch_bounds__(kBlock, MAXP <= 2 ? 8 : MAXP <= 6 ? 5 : 3)
poker_packed_reset_kernel(PulsePokerPacked S, PokerParams pr, const int32_t* __restrict__ prefixed, uint64_t seed,
                          uint64_t env_base, uint32_t hand, int rotation, int first_reset, int starting_bbs,
                          int max_bbs, float* __restrict__ obs_out, long long n) {
  extern __shared__ float4 smem4[];
  if (MT > 0) pr.M = MT;
  constexpr int WT = MT > 0 ? 13 + 3 * (MT - 1) : 13 + 3 * 15;
  const int W = 13 + 3 * (pr.M - 1);
  PackedDeckShared dk{reinterpret_cast<uint8_t*>(smem4) + 4 * threadIdx.x, 4 * kBlock};
  float4* stage = smem4 + kDeckSmemBytes / 16;
  for (long long base = (long long)blockIdx.x * kBlock; base < n; base += (long long)gridDim.x * kBlock) {
    const long long g = base + threadIdx.x;
    float row[WT];
    if (g < n)
      packed_reset_one<MAXP>(S, pr, prefixed, seed, env_base, hand, rotation, first_reset, starting_bbs, max_bbs,
                             obs_out ? row : nullptr, g, n, dk);
    if (obs_out) stage_obs_rows<MT, WT>(row, stage, obs_out, base, g, n, W);  // obs_out is uniform: barriers are safe
  }
}

// Redeal pass of auto_reset.  Marked tables are usually a small, scattered fraction of the batch, and a deal is
// ~1,500 instructions: one thread per table would leave most lanes of every warp idle through the whole deal
// (measured: 3 of 32 lanes active with 10% of the tables marked).  So each block first compacts the marked
// tables of its tile (kRedealTile tables, 4 B coalesced read of hand_ctr each) into a list in shared memory, then
// its threads deal list entries densely.  Dealt tables get their observation row rewritten.
constexpr int kRedealTile = 8 * kBlock;

template <int MAXP>
__global__ void __launch_bounds__(kBlock, MAXP <= 2 ? 8 : MAXP <= 6 ? 5 : 3)
poker_packed_redeal_kernel(PulsePokerPacked S, PokerParams pr, uint64_t seed, uint64_t env_base, int starting_bbs,
                           int max_bbs, float* __restrict__ obs_out, long long n) {
  extern __shared__ float4 smem4[];
  __shared__ int list[kRedealTile];
  __shared__ int count;
  const int W = 13 + 3 * (pr.M - 1);
  const bool vec = (W & 3) == 0 && (reinterpret_cast<uintptr_t>(obs_out) & 15u) == 0;
  PackedDeckShared dk{reinterpret_cast<uint8_t*>(smem4) + 4 * threadIdx.x, 4 * kBlock};
  for (long long base = (long long)blockIdx.x * kRedealTile; base < n; base += (long long)gridDim.x * kRedealTile) {
    if (threadIdx.x == 0) count = 0;
    __syncthreads();
    // all eight counter words are requested before any is looked at (as one load-test-append per iteration the
    // loads went out one DRAM round trip after another: two thirds of the kernel's stall samples, r1r capture);
    // a warp appends its marked tables with one shared atomic
    uint32_t ctr[kRedealTile / kBlock];
#pragma unroll
    for (int k = 0; k < kRedealTile / kBlock; ++k) {
      const long long g = base + k * kBlock + threadIdx.x;
      ctr[k] = g < n ? S.hand_ctr[g] : 0u;
    }
    const unsigned lane = threadIdx.x & 31u;
#pragma unroll
    for (int k = 0; k < kRedealTile / kBlock; ++k) {
      const bool marked = (ctr[k] & PKP_REDEAL_MARK) != 0;
      const unsigned bal = __ballot_sync(0xFFFFFFFFu, marked);
      if (bal) {
        int pos = 0;
        if (lane == 0) pos = atomicAdd(&count, __popc(bal));
        pos = __shfl_sync(0xFFFFFFFFu, pos, 0);
        if (marked) list[pos + __popc(bal & ((1u << lane) - 1u))] = k * kBlock + (int)threadIdx.x;
      }
    }
    __syncthreads();
    const int cnt = count;
    for (int i = threadIdx.x; i < cnt; i += kBlock) {
      const long long g = base + list[i];
      float row[13 + 3 * 15];
      packed_redeal_one<MAXP>(S, pr, seed, env_base, starting_bbs, max_bbs, row, g, n, dk);
      float* dst = obs_out + g * W;
      if (vec) {
        for (int c = 0; c < W; c += 4)
          *reinterpret_cast<float4*>(dst + c) = make_float4(row[c], row[c + 1], row[c + 2], row[c + 3]);
      } else {
        for (int c = 0; c < W; ++c) dst[c] = row[c];
      }
    }
    __syncthreads();
  }
}

// K steps per launch with every seat played on the device (SURVEY.md 8(b) pulse_poker_rollout): at the reference's own
// benchmark size (N = 256, benchmarking/Poker) one step is a few microseconds of work behind a launch, so the loop over
// steps runs inside the kernel.  A thread owns its table for all K steps; the table's words go through global memory
// between steps (they stay in L1/L2), so the per-step body is exactly the single-step one.  Outputs step-major.
template <int MAXP>
__global__ void __launch_bounds__(kBlock)
poker_packed_rollout_kernel(PulsePokerPacked S, PokerParams pr, PackedPolicy pol, int K, int auto_reset, uint64_t seed,
                            uint64_t env_base, int starting_bbs, int max_bbs, long long n, float* __restrict__ obs_final,
                            float* __restrict__ reward_out, uint8_t* __restrict__ done_out, int64_t* __restrict__ action_out,
                            const uint8_t* __restrict__ tape) {
  const int W = 13 + 3 * (pr.M - 1);
  if (auto_reset == 0) {
    // No redeals inside the launch: the table is loaded once, stepped K times in registers and stored once -- the K - 1
    // intermediate load / decode / pack / store rounds (a fifth of the single step's instructions and all of its state
    // traffic) never happen.  What a step writes: reward, done and, when asked, the action played.
    for (long long g = (long long)blockIdx.x * kBlock + threadIdx.x; g < n; g += (long long)gridDim.x * kBlock) {
      PackedEv<MAXP, (MAXP <= 2)> evw{S.ev, g, n};
      evw.prefetch(pr.AP);
      PokerTable<MAXP> t;
      uint32_t fb;
      packed_load(t, fb, S, g, n, pr.AP);
      const PokerTable<MAXP> t0 = t;
      int prev_inv = 0;
      for (int k = 0; k < K; ++k) {
        long long a;
        if (tape) {                                       // supplied actions, uint8 [K, N]; 255 = "negative" (matches no branch)
          a = (long long)tape[(long long)k * n + g];
          if (a == 255) a = -1;
        } else {
          const int kind = pol.kinds.kind[t.idx & 15];    // build_actions: from the pre-step state, done or not (utils.py:108-123)
          const PulseU4 b = pulse_tape(seed, env_base + (uint64_t)g, PULSE_STREAM_ACTION, pol.counter + (uint32_t)k);
          a = pulse_policy_action(kind, pk_sel(t.hole0, t.idx), pk_sel(t.hole1, t.idx), (float)t.pot, b.x, b.y);
        }
        if (action_out) action_out[(long long)k * n + g] = a;
        const PackedStepResult r = packed_step_regs<MAXP>(t, fb, evw, pr, a, prev_inv);
        reward_out[(long long)k * n + g] = r.reward;
        done_out[(long long)k * n + g] = r.done;
      }
      if (S.prev_invested) S.prev_invested[g] = prev_inv;
      packed_store(t, t0, S, g, n, pr.AP);
      if (S.idx && t.idx != t0.idx) S.idx[g] = t.idx;
      if (obs_final) {
        float row[13 + 3 * 15];
        pk_obs(t, pr, row);
        for (int c = 0; c < W; ++c) obs_final[g * W + c] = row[c];
      }
    }
    return;
  }
  for (long long g = (long long)blockIdx.x * kBlock + threadIdx.x; g < n; g += (long long)gridDim.x * kBlock) {
    PackedPolicy p = pol;
    for (int k = 0; k < K; ++k) {
      p.counter = pol.counter + (uint32_t)k;
      p.chosen = action_out ? action_out + (long long)k * n : nullptr;
      PokerTable<MAXP> t;
      const PackedStepResult r = packed_step_core<MAXP, true, 1>(S, pr, static_cast<const int64_t*>(nullptr), auto_reset & 1, seed,
                                                                 env_base, starting_bbs, max_bbs, g, n, &p, t);
      reward_out[(long long)k * n + g] = r.reward;
      done_out[(long long)k * n + g] = r.done;
      if (k == K - 1 && obs_final) {
        float row[13 + 3 * 15];
        pk_obs(t, pr, row);
        for (int c = 0; c < W; ++c) obs_final[g * W + c] = row[c];
      }
    }
  }
}

template <int MAXP>
__global__ void __launch_bounds__(kBlock)
poker_packed_unpack_kernel(PulsePokerPacked S, PulsePokerState C, PokerParams pr, long long n, bool with_decks,
                           const int32_t* __restrict__ prefixed, uint64_t seed, uint64_t env_base, bool fresh) {
  for (long long g = (long long)blockIdx.x * kBlock + threadIdx.x; g < n; g += (long long)gridDim.x * kBlock)
    packed_unpack_one<MAXP>(S, C, pr, g, n, with_decks, prefixed, seed, env_base, fresh);
}

template <int MAXP>
__global__ void __launch_bounds__(kBlock)
poker_packed_pack_kernel(PulsePokerPacked S, PulsePokerState C, PokerParams pr, long long n, int* __restrict__ bad) {
  int mine = 0;
  for (long long g = (long long)blockIdx.x * kBlock + threadIdx.x; g < n; g += (long long)gridDim.x * kBlock)
    mine += packed_pack_one<MAXP>(S, C, pr, g, n) ? 0 : 1;
  if (bad) {
    mine = __reduce_add_sync(0xFFFFFFFFu, mine);
    if ((threadIdx.x & 31) == 0 && mine) atomicAdd(bad, mine);
  }
}

// one thread per (table, seat): seat-major so a warp reads 32 consecutive tables' hole cards and table words
__global__ void __launch_bounds__(kBlock)
poker_packed_strength_kernel(PulsePokerPacked S, int AP, float* __restrict__ out, long long n) {
  const long long total = n * AP;
  for (long long i = (long long)blockIdx.x * kBlock + threadIdx.x; i < total; i += (long long)gridDim.x * kBlock) {
    const int s = (int)(i / n);
    const long long g = i - (long long)s * n;
    const int stage = (int)(S.tbl[4 * g] & 7u);
    const uint32_t fb = S.tbl2[2 * g + 1];
    const uint32_t h = S.hole[(long long)s * n + g];
    out[g * AP + s] = packed_enumerated_strength(stage, (int)(h & 63u), (int)((h >> 6) & 63u), fb);
  }
}

// Showdown equity: one WARP per table -- tables of a batch sit on different streets (1 to 1,081 boards each), so
// lanes split a table's boards instead of tables splitting a warp.  Every lane loads the table (same addresses: one
// transaction per warp), scores its boards, and the per-seat integer sums are added across the warp.
template <int MAXP>
__global__ void __launch_bounds__(kBlock)
poker_packed_showdown_equity_kernel(PulsePokerPacked S, PokerParams pr, int samples, uint64_t seed, uint64_t env_base,
                                    uint32_t counter, float* __restrict__ out, long long n) {
  const int lane = threadIdx.x & 31;
  const long long nw = ((long long)gridDim.x * kBlock) >> 5;
  for (long long g = ((long long)blockIdx.x * kBlock + threadIdx.x) >> 5; g < n; g += nw) {
    PokerTable<MAXP> t;
    uint32_t fb;
    packed_load(t, fb, S, g, n, pr.AP);
    int32_t num[MAXP];
#pragma unroll
    for (int s = 0; s < MAXP; ++s) num[s] = 0;
    const int boards = packed_showdown_equity_part<MAXP>(t, pr.AP, samples, seed, env_base + (uint64_t)g, counter, lane,
                                                         32, num);
    float mine = 0.0f;
#pragma unroll
    for (int s = 0; s < MAXP; ++s) {
      int v = num[s];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, o);
      if (lane == s) mine = pkq_equity((int64_t)v, boards);
    }
    if (lane < pr.AP) out[g * pr.AP + lane] = mine;
  }
}

int check_cfg(const void* st, const PulsePokerConfig* cfg) {
  if (!st || !cfg) { pulse_set_error("null state/config"); return PULSE_EINVAL; }
  if (cfg->n_players < 2 || cfg->n_players > 10) { pulse_set_error("n_players must be in 2..10"); return PULSE_EINVAL; }
  if (cfg->active_players < 2 || cfg->active_players > cfg->n_players) { pulse_set_error("active_players must be in 2..n_players"); return PULSE_EINVAL; }
  if (cfg->max_players < cfg->active_players || cfg->max_players > 16) { pulse_set_error("max_players must be in active_players..16"); return PULSE_EINVAL; }
  if (cfg->n_games < 0) { pulse_set_error("n_games < 0"); return PULSE_EINVAL; }
  return PULSE_OK;
}

PokerParams make_params(const PulsePokerConfig* c) {
  PokerParams p; p.w1 = c->w1; p.w2 = c->w2; p.K = c->K; p.alpha = c->alpha;
  p.P = c->n_players; p.AP = c->active_players; p.M = c->max_players;
  return p;
}

}  // namespace

// how many CTAs ahead the step kernel prefetches into L2: one resident wave (148 SMs x resident CTAs per SM);
// PULSE_POKER_L2_AHEAD overrides (0 turns the prefetch off) for experiments
static int l2_ahead_for(int minb, int max_players) {
  static const int env = getenv("PULSE_POKER_L2_AHEAD") ? atoi(getenv("PULSE_POKER_L2_AHEAD")) : -1;
  static const int env_ev = getenv("PULSE_POKER_L2_EV") ? atoi(getenv("PULSE_POKER_L2_EV")) : -1;
  // measured on B200 (profiles/r1r_l2_ahead.md): two CTAs per SM ahead heads-up, one with more seats; further ahead
  // the streamed observation writes push the prefetched lines out of L2 again before they are used
  const int ahead = env >= 0 ? env : (max_players <= 2 ? 296 : 148);
  (void)minb;
  // heads-up reads both seats' hand-value words with the rest of the table, so they belong to the prefetch set;
  // with more seats only the actor's (or the showdown's) words are read
  const int ev = env_ev >= 0 ? env_ev : (max_players <= 2);
  return ahead > 0 ? (ahead | (ev ? 1 << 30 : 0)) : 0;
}

static int launch_redeal(const PulsePokerPacked* st, const PulsePokerConfig* cfg, const PokerParams& pr, uint64_t seed,
                         uint64_t env_base, float* obs_out, cudaStream_t stream);

#define PKP_DISPATCH(P, CALL)                            \
  do {                                                   \
    if ((P) <= 2) { constexpr int MP = 2; CALL; }        \
    else if ((P) <= 4) { constexpr int MP = 4; CALL; }   \
    else if ((P) <= 6) { constexpr int MP = 6; CALL; }   \
    else { constexpr int MP = 10; CALL; }                \
  } while (0)

static int launch_redeal(const PulsePokerPacked* st, const PulsePokerConfig* cfg, const PokerParams& pr, uint64_t seed,
                         uint64_t env_base, float* obs_out, cudaStream_t stream) {
  const int grid = pulse_grid_for(((long long)cfg->n_games + 7) / 8, kBlock);   // one block per kRedealTile tables
  PKP_DISPATCH(pr.P, (poker_packed_redeal_kernel<MP><<<grid, kBlock, kDeckSmemBytes, stream>>>(
                         *st, pr, seed, env_base, cfg->starting_bbs, cfg->max_bbs, obs_out, (long long)cfg->n_games)));
  return pulse_check_launch("pulse_poker_packed_step (redeal pass)");
}

extern "C" {

int pulse_poker_packed_reset(const PulsePokerPacked* st, const PulsePokerConfig* cfg, const int32_t* prefixed,
                             uint64_t seed, uint64_t env_base, uint32_t hand_counter, int rotation, int first_reset,
                             float* obs_out, cudaStream_t stream) {
  int rc = check_cfg(st, cfg);
  if (rc) return rc;
  if (cfg->n_games == 0) return PULSE_OK;
  PokerParams pr = make_params(cfg);
  const int grid = pulse_grid_for(cfg->n_games, kBlock);
  const int W = 13 + 3 * (pr.M - 1);
#define PKP_RESET(MP, MT, STAGE)                                                                                   \
  poker_packed_reset_kernel<MP, MT><<<grid, kBlock, kDeckSmemBytes + (STAGE), stream>>>(                            \
      *st, pr, prefixed, seed, env_base, hand_counter, rotation, first_reset, cfg->starting_bbs, cfg->max_bbs,     \
      obs_out, (long long)cfg->n_games)
#define PKP_SMEM4(MT) ((size_t)kBlock * (((13 + 3 * ((MT) - 1)) / 4) | 1) * sizeof(float4))
  const bool aligned = (reinterpret_cast<uintptr_t>(obs_out) & 15u) == 0;
  if (pr.P <= 2 && pr.M == 2 && aligned) PKP_RESET(2, 2, PKP_SMEM4(2));
  else if (pr.P > 4 && pr.P <= 6 && pr.M == 6 && aligned) PKP_RESET(6, 6, PKP_SMEM4(6));
  else if (pr.P > 6 && pr.M == 10 && aligned) PKP_RESET(10, 10, PKP_SMEM4(10));
  else {
    const size_t stage = (size_t)kBlock * (W + 1) * sizeof(float);
    PKP_DISPATCH(pr.P, PKP_RESET(MP, 0, stage));
  }
#undef PKP_RESET
#undef PKP_SMEM4
  return pulse_check_launch("pulse_poker_packed_reset");
}

int pulse_poker_packed_step(const PulsePokerPacked* st, const PulsePokerConfig* cfg, const int64_t* actions,
                            float* obs_out, float* reward_out, uint8_t* done_out, int auto_reset, uint64_t seed,
                            uint64_t env_base, cudaStream_t stream) {
  int rc = check_cfg(st, cfg);
  if (rc) return rc;
  if (cfg->n_games == 0) return PULSE_OK;
  PULSE_REQUIRE(actions && obs_out && reward_out && done_out, "null buffer");
  PULSE_REQUIRE(!auto_reset || st->hand_ctr, "auto_reset needs hand_ctr");
  PokerParams pr = make_params(cfg);
  const int W = 13 + 3 * (pr.M - 1);
  const int grid = pulse_grid_for(cfg->n_games, kBlock);
  // heads-up rows: direct 256-bit stores (PULSE_POKER_OBS_DIRECT=0 keeps the staged write-out, for A/B runs)
  static const bool direct_env = !(getenv("PULSE_POKER_OBS_DIRECT") && atoi(getenv("PULSE_POKER_OBS_DIRECT")) == 0);
  const bool direct = direct_env && (reinterpret_cast<uintptr_t>(obs_out) & 31u) == 0;
  // auto_reset: finished tables are dealt by the compacting redeal pass that follows the step kernel.  Dealing them
  // inside the step kernel, tile by tile, was built and measured in round 2 (profiles/r2e_bench_lines.jsonl): 4-15%
  // slower on every auto-reset workload -- a 128-table tile yields ~19 deals, two thirds of one warp, and the CTA waits
  // for that warp -- and was removed again.
#define PKP_STEP_A(MP, MT, MINB, AUTO, SMEM)                                                              \
  do {                                                                                                    \
    if ((MT) == 2 && direct)                                                                              \
      poker_packed_step_kernel<MP, MT, MINB, false, AUTO, true><<<grid, kBlock, 0, stream>>>(              \
          *st, pr, actions, obs_out, reward_out, done_out, AUTO, seed, env_base, cfg->starting_bbs, cfg->max_bbs, \
          (long long)cfg->n_games, l2_ahead_for(MINB, MP));                                               \
    else                                                                                                  \
      poker_packed_step_kernel<MP, MT, MINB, false, AUTO><<<grid, kBlock, SMEM, stream>>>(                 \
          *st, pr, actions, obs_out, reward_out, done_out, AUTO, seed, env_base, cfg->starting_bbs, cfg->max_bbs, \
          (long long)cfg->n_games, l2_ahead_for(MINB, MP));                                               \
  } while (0)
#define PKP_STEP_B(MP, MT, MINB, SMEM)                                  \
  do {                                                                  \
    if (auto_reset) PKP_STEP_A(MP, MT, MINB, 2, SMEM);                  \
    else PKP_STEP_A(MP, MT, MINB, 0, SMEM);                             \
  } while (0)
  // resident CTAs per SM the step kernel is compiled for: 10 (48 registers) heads-up, 6 / 4 for the six- / ten-seat
  // builds (below), 8 (64) for the generic ones -- measured on B200 (profiles/r1c_poker_launch_bounds.md, r1r_exp5/6);
  // PULSE_POKER_MINB overrides for experiments
  static const int minb_env = getenv("PULSE_POKER_MINB") ? atoi(getenv("PULSE_POKER_MINB")) : 0;
#define PKP_STEP(MP, MT, SMEM)                                    \
  do {                                                            \
    const int minb = minb_env > 0 ? minb_env : ((MP) <= 2 && !direct ? 10 : 8); \
    if (minb >= 12) PKP_STEP_B(MP, MT, 12, SMEM);                 \
    else if (minb >= 10) PKP_STEP_B(MP, MT, 10, SMEM);            \
    else PKP_STEP_B(MP, MT, 8, SMEM);                             \
  } while (0)
#define PKP_SMEM4(MT) ((size_t)kBlock * (((13 + 3 * ((MT) - 1)) / 4) | 1) * sizeof(float4))
  const bool aligned = (reinterpret_cast<uintptr_t>(obs_out) & 15u) == 0;
  if (pr.P <= 2 && pr.M == 2 && aligned) PKP_STEP(2, 2, PKP_SMEM4(2));
  // More seats: the table no longer fits the 64 registers of 8 CTAs/SM, and with the L2 look-ahead hiding most of the
  // load latency, fewer and fatter CTAs win.  Measured on B200 (profiles/r1r_exp5_minb_lines.jsonl, r1r_exp6_...):
  //   six seats: 6 CTAs/SM (80 registers) 0.83 of HBM, 8 (64) 0.80, 5 (96) 0.80, 10 (48) 0.65, 12 (40) 0.54
  //   ten seats: 4 CTAs/SM (128 registers) 0.67, 5 (96) 0.62, 3 (160) 0.59, 8 (64, 320 B spill) 0.49, 10 (48) 0.36
  else if (pr.P > 4 && pr.P <= 6 && pr.M == 6 && aligned) {
    if (minb_env == 5) PKP_STEP_B(6, 6, 5, PKP_SMEM4(6));
    else if (minb_env >= 8) PKP_STEP(6, 6, PKP_SMEM4(6));
    else PKP_STEP_B(6, 6, 6, PKP_SMEM4(6));
  }
  else if (pr.P > 6 && pr.M == 10 && aligned) {
    if (minb_env == 3) PKP_STEP_B(10, 10, 3, PKP_SMEM4(10));
    else if (minb_env == 5) PKP_STEP_B(10, 10, 5, PKP_SMEM4(10));
    else if (minb_env >= 8) PKP_STEP(10, 10, PKP_SMEM4(10));
    else PKP_STEP_B(10, 10, 4, PKP_SMEM4(10));
  }
  else {
    const size_t smem = (size_t)kBlock * (W + 1) * sizeof(float);
    PKP_DISPATCH(pr.P, PKP_STEP(MP, 0, smem));
  }
#undef PKP_STEP
#undef PKP_STEP_B
#undef PKP_STEP_A
#undef PKP_SMEM4
  rc = pulse_check_launch("pulse_poker_packed_step");
  if (rc || !auto_reset) return rc;
  return launch_redeal(st, cfg, pr, seed, env_base, obs_out, stream);
}

int pulse_poker_packed_step_policy(const PulsePokerPacked* st, const PulsePokerConfig* cfg,
                                   const int64_t* actions_or_null, const uint8_t* seat_kinds,
                                   const uint32_t* chance2_or_null, uint32_t counter, int64_t* chosen_out_or_null,
                                   int32_t* seat_idx_out_or_null, float* obs_out, float* reward_out, uint8_t* done_out, int auto_reset,
                                   uint64_t seed, uint64_t env_base, cudaStream_t stream) {
  int rc = check_cfg(st, cfg);
  if (rc) return rc;
  PULSE_REQUIRE(seat_kinds != nullptr, "seat_kinds is null");
  PackedPolicy pol;
  bool any_external = false;
  for (int s = 0; s < 16; ++s) {
    pol.kinds.kind[s] = s < cfg->n_players ? seat_kinds[s] : (uint8_t)PULSE_AGENT_EXTERNAL;
    if (s < cfg->n_players) {
      PULSE_REQUIRE(pol.kinds.kind[s] <= PULSE_AGENT_SMALL_BALL, "unknown agent kind");
      any_external |= pol.kinds.kind[s] == PULSE_AGENT_EXTERNAL;
    }
  }
  PULSE_REQUIRE(!any_external || actions_or_null, "actions are required when a seat is PULSE_AGENT_EXTERNAL");
  if (cfg->n_games == 0) return PULSE_OK;
  PULSE_REQUIRE(obs_out && reward_out && done_out, "null buffer");
  PULSE_REQUIRE(!auto_reset || st->hand_ctr, "auto_reset needs hand_ctr");
  pol.chance2 = chance2_or_null; pol.chosen = chosen_out_or_null; pol.seat_before = seat_idx_out_or_null;
  pol.counter = counter;
  PokerParams pr = make_params(cfg);
  const int W = 13 + 3 * (pr.M - 1);
  const int grid = pulse_grid_for(cfg->n_games, kBlock);
  static const bool direct_env = !(getenv("PULSE_POKER_OBS_DIRECT") && atoi(getenv("PULSE_POKER_OBS_DIRECT")) == 0);
  const bool direct = direct_env && (reinterpret_cast<uintptr_t>(obs_out) & 31u) == 0;
#define PKP_POL_A(MP, MT, MINB, AUTO, SMEM)                                                                \
  do {                                                                                                     \
    if ((MT) == 2 && direct)                                                                               \
      poker_packed_step_kernel<MP, MT, MINB, true, AUTO, true><<<grid, kBlock, 0, stream>>>(                \
          *st, pr, actions_or_null, obs_out, reward_out, done_out, AUTO, seed, env_base, cfg->starting_bbs,  \
          cfg->max_bbs, (long long)cfg->n_games, l2_ahead_for(MINB, MP), pol);                             \
    else                                                                                                   \
      poker_packed_step_kernel<MP, MT, MINB, true, AUTO><<<grid, kBlock, SMEM, stream>>>(                   \
          *st, pr, actions_or_null, obs_out, reward_out, done_out, AUTO, seed, env_base, cfg->starting_bbs,  \
          cfg->max_bbs, (long long)cfg->n_games, l2_ahead_for(MINB, MP), pol);                             \
  } while (0)
#define PKP_POL(MP, MT, MINB, SMEM)                                     \
  do {                                                                  \
    if (auto_reset) PKP_POL_A(MP, MT, MINB, 2, SMEM);                   \
    else PKP_POL_A(MP, MT, MINB, 0, SMEM);                              \
  } while (0)
#define PKP_SMEM4(MT) ((size_t)kBlock * (((13 + 3 * ((MT) - 1)) / 4) | 1) * sizeof(float4))
  const bool aligned = (reinterpret_cast<uintptr_t>(obs_out) & 15u) == 0;
  static const int pol_minb = getenv("PULSE_POKER_POLICY_MINB") ? atoi(getenv("PULSE_POKER_POLICY_MINB")) : 8;  // 63 registers: measured 0.443 vs 0.477 ms per train-loop step
  if (pr.P <= 2 && pr.M == 2 && aligned) {
    if (pol_minb >= 10) PKP_POL(2, 2, 10, PKP_SMEM4(2));
    else PKP_POL(2, 2, 8, PKP_SMEM4(2));
  }
  else if (pr.P > 4 && pr.P <= 6 && pr.M == 6 && aligned) PKP_POL(6, 6, 8, PKP_SMEM4(6));
  else if (pr.P > 6 && pr.M == 10 && aligned) PKP_POL(10, 10, 8, PKP_SMEM4(10));
  else {
    const size_t smem = (size_t)kBlock * (W + 1) * sizeof(float);
    PKP_DISPATCH(pr.P, PKP_POL(MP, 0, 8, smem));
  }
#undef PKP_POL
#undef PKP_POL_A
#undef PKP_SMEM4
  rc = pulse_check_launch("pulse_poker_packed_step_policy");
  if (rc || !auto_reset) return rc;
  return launch_redeal(st, cfg, pr, seed, env_base, obs_out, stream);
}

int pulse_poker_packed_rollout(const PulsePokerPacked* st, const PulsePokerConfig* cfg, const uint8_t* seat_kinds,
                               uint32_t counter, int n_steps, int auto_reset, uint64_t seed, uint64_t env_base,
                               float* obs_final_or_null, float* reward_out, uint8_t* done_out, int64_t* actions_out_or_null,
                               cudaStream_t stream) {
  int rc = check_cfg(st, cfg);
  if (rc) return rc;
  PULSE_REQUIRE(seat_kinds != nullptr, "seat_kinds is null");
  PULSE_REQUIRE(n_steps >= 0, "n_steps < 0");
  PackedPolicy pol;
  for (int s = 0; s < 16; ++s) {
    pol.kinds.kind[s] = s < cfg->n_players ? seat_kinds[s] : (uint8_t)PULSE_AGENT_RANDOM;
    if (s < cfg->n_players)
      PULSE_REQUIRE(pol.kinds.kind[s] >= PULSE_AGENT_RANDOM && pol.kinds.kind[s] <= PULSE_AGENT_SMALL_BALL,
                    "a rollout plays every seat on the device: no PULSE_AGENT_EXTERNAL seats");
  }
  if (cfg->n_games == 0 || n_steps == 0) return PULSE_OK;
  PULSE_REQUIRE(reward_out && done_out, "null buffer");
  PULSE_REQUIRE(!auto_reset || st->hand_ctr, "auto_reset needs hand_ctr");
  pol.chance2 = nullptr; pol.chosen = nullptr; pol.seat_before = nullptr; pol.counter = counter;
  PokerParams pr = make_params(cfg);
  const int grid = pulse_grid_for(cfg->n_games, kBlock);
  // PULSE_POKER_ROLLOUT_REGS=0: step through global memory between steps even without auto_reset (the A/B switch of
  // the registers-resident path; bit 1 of the kernel's auto_reset argument)
  static const bool regs_env = !(getenv("PULSE_POKER_ROLLOUT_REGS") && atoi(getenv("PULSE_POKER_ROLLOUT_REGS")) == 0);
  if (!auto_reset && !regs_env) auto_reset = 2;
  else auto_reset = auto_reset ? 1 : 0;
  PKP_DISPATCH(pr.P, (poker_packed_rollout_kernel<MP><<<grid, kBlock, 0, stream>>>(
                         *st, pr, pol, n_steps, auto_reset, seed, env_base, cfg->starting_bbs, cfg->max_bbs,
                         (long long)cfg->n_games, obs_final_or_null, reward_out, done_out, actions_out_or_null, nullptr)));
  return pulse_check_launch("pulse_poker_packed_rollout");
}

int pulse_poker_packed_rollout_actions(const PulsePokerPacked* st, const PulsePokerConfig* cfg, const uint8_t* actions,
                                       int n_steps, float* obs_final_or_null, float* reward_out, uint8_t* done_out,
                                       cudaStream_t stream) {
  int rc = check_cfg(st, cfg);
  if (rc) return rc;
  PULSE_REQUIRE(n_steps >= 0, "n_steps < 0");
  if (cfg->n_games == 0 || n_steps == 0) return PULSE_OK;
  PULSE_REQUIRE(actions && reward_out && done_out, "null buffer");
  PackedPolicy pol;
  for (int s = 0; s < 16; ++s) pol.kinds.kind[s] = (uint8_t)PULSE_AGENT_EXTERNAL;
  pol.chance2 = nullptr; pol.chosen = nullptr; pol.seat_before = nullptr; pol.counter = 0;
  PokerParams pr = make_params(cfg);
  const int grid = pulse_grid_for(cfg->n_games, kBlock);
  PKP_DISPATCH(pr.P, (poker_packed_rollout_kernel<MP><<<grid, kBlock, 0, stream>>>(
                         *st, pr, pol, n_steps, 0, 0, 0, cfg->starting_bbs, cfg->max_bbs, (long long)cfg->n_games,
                         obs_final_or_null, reward_out, done_out, nullptr, actions)));
  return pulse_check_launch("pulse_poker_packed_rollout_actions");
}

int pulse_poker_packed_showdown_equity(const PulsePokerPacked* st, const PulsePokerConfig* cfg, float* equity_out,
                                       int preflop_samples, uint64_t seed, uint64_t env_base, uint32_t counter,
                                       cudaStream_t stream) {
  int rc = check_cfg(st, cfg);
  if (rc) return rc;
  PULSE_REQUIRE(preflop_samples >= 1 && preflop_samples <= PKQ_MAX_SAMPLES, "preflop_samples must be in 1..65536");
  if (cfg->n_games == 0) return PULSE_OK;
  PULSE_REQUIRE(equity_out != nullptr, "equity_out is null");
  PokerParams pr = make_params(cfg);
  const int grid = pulse_grid_for((long long)cfg->n_games * 32, kBlock);
  PKP_DISPATCH(pr.P, (poker_packed_showdown_equity_kernel<MP><<<grid, kBlock, 0, stream>>>(
                         *st, pr, preflop_samples, seed, env_base, counter, equity_out, (long long)cfg->n_games)));
  return pulse_check_launch("pulse_poker_packed_showdown_equity");
}

int pulse_poker_pack(const PulsePokerState* compat, const PulsePokerPacked* st, const PulsePokerConfig* cfg,
                     int* bad_count_or_null, cudaStream_t stream) {
  int rc = check_cfg(st, cfg);
  if (rc) return rc;
  PULSE_REQUIRE(compat != nullptr, "compat state is null");
  if (cfg->n_games == 0) return PULSE_OK;
  PULSE_REQUIRE(compat->decks && compat->deck_positions && compat->board && compat->pots && compat->stages && compat->stacks &&
                    compat->hands && compat->current_round_bet && compat->total_invested && compat->status && compat->button &&
                    compat->idx && compat->highest && compat->agg && compat->acted && compat->last_raise_size && compat->is_done,
                "pack needs every state tensor (equities, equity_dirty, prev_invested, sb, bb are optional)");
  PokerParams pr = make_params(cfg);
  const int grid = pulse_grid_for(cfg->n_games, kBlock);
  PKP_DISPATCH(pr.P, (poker_packed_pack_kernel<MP><<<grid, kBlock, 0, stream>>>(*st, *compat, pr, (long long)cfg->n_games,
                                                                               bad_count_or_null)));
  return pulse_check_launch("pulse_poker_pack");
}

int pulse_poker_packed_enumerated_strength(const PulsePokerPacked* st, const PulsePokerConfig* cfg, float* strength_out,
                                           cudaStream_t stream) {
  int rc = check_cfg(st, cfg);
  if (rc) return rc;
  if (cfg->n_games == 0) return PULSE_OK;
  PULSE_REQUIRE(strength_out != nullptr, "strength_out is null");
  const long long total = (long long)cfg->n_games * cfg->active_players;
  poker_packed_strength_kernel<<<pulse_grid_for(total, kBlock), kBlock, 0, stream>>>(*st, cfg->active_players,
                                                                                    strength_out, (long long)cfg->n_games);
  return pulse_check_launch("pulse_poker_packed_enumerated_strength");
}

static int launch_unpack(const PulsePokerPacked* st, const PulsePokerState* compat, const PulsePokerConfig* cfg,
                         bool with_decks, const int32_t* prefixed, uint64_t seed, uint64_t env_base, bool fresh,
                         cudaStream_t stream, const char* what) {
  int rc = check_cfg(st, cfg);
  if (rc) return rc;
  PULSE_REQUIRE(compat != nullptr, "compat state is null");
  PULSE_REQUIRE(!compat->equities || st->ev, "equities need the packed hand-value words");
  if (cfg->n_games == 0) return PULSE_OK;
  PokerParams pr = make_params(cfg);
  const int grid = pulse_grid_for(cfg->n_games, kBlock);
  PKP_DISPATCH(pr.P, (poker_packed_unpack_kernel<MP><<<grid, kBlock, 0, stream>>>(
                         *st, *compat, pr, (long long)cfg->n_games, with_decks, prefixed, seed, env_base, fresh)));
  return pulse_check_launch(what);
}

int pulse_poker_unpack(const PulsePokerPacked* st, const PulsePokerState* compat, const PulsePokerConfig* cfg,
                       cudaStream_t stream) {
  return launch_unpack(st, compat, cfg, false, nullptr, 0, 0, false, stream, "pulse_poker_unpack");
}

int pulse_poker_unpack_full(const PulsePokerPacked* st, const PulsePokerState* compat, const PulsePokerConfig* cfg,
                            const int32_t* prefixed_decks_or_null, uint64_t seed, uint64_t env_base, int fresh,
                            cudaStream_t stream) {
  return launch_unpack(st, compat, cfg, true, prefixed_decks_or_null, seed, env_base, fresh != 0, stream,
                       "pulse_poker_unpack_full");
}
}
