// 2048 / TFE kernels and their C ABI (include/pulse_envs.h).  Board rules and the packed layout are in
// tfe_body.cuh.  One thread per board, one launch per step: 8 B board in, 8 B out, optional int32
// side x side observation staged through shared memory so the [N, side*side] slab is written with
// coalesced stores.  HBM-bound (106 B per board-step with the int32 observation, 42 B
// without).
#include "tfe_body.cuh"
#include <stdlib.h>

namespace {
constexpr int kBlock = 256;

template <int SIDE>
__device__ __forceinline__ void tfe_write_obs(int32_t* smem, const uint64_t b, const bool valid, const long long base,
                                              const long long n, int32_t* __restrict__ obs_out) {
  constexpr int W = SIDE * SIDE, S = W + 1;  // odd row stride: conflict-free both ways
  if (valid) {
    int32_t row[W];
    tfe_unpack_board<SIDE>(b, row);
#pragma unroll
    for (int i = 0; i < W; ++i) smem[threadIdx.x * S + i] = row[i];
  }
  __syncthreads();
  const long long rows = (n - base) < kBlock ? (n - base) : kBlock;
  const int total = (int)rows * W;
  int32_t* dst = obs_out + base * W;
#pragma unroll
  for (int it = 0; it < W; ++it) {
    const int q = threadIdx.x + it * kBlock;
    if (q < total) dst[q] = smem[(q / W) * S + (q % W)];
  }
  __syncthreads();
}

// ---- step kernel ------------------------------------------------------------------------------------
// Persistent: one 1024-thread CTA per SM (grid <= 148).  Each CTA stages the 65,536-entry row-move table
// (uint16 squashed row + uint8 score code = 192 KB) in shared memory once by TMA bulk copy
// (cp.async.bulk + mbarrier, overlapped with the first board's loads and Philox draw), then its threads walk the boards
// with a grid stride, prefetching the next board while the current one is processed.  A move is four
// table lookups instead of four slide-and-merge computations.  Observations leave through a per-warp
// staging slot: the warp's 32 packed boards go to shared memory and come back out as the warp's contiguous
// [32, side*side] int32 slab in fully coalesced 16 B stores (side 4) -- no CTA-wide barrier in the loop.
constexpr int kStepBlock = 1024;
constexpr int kTableBytes = 65536 * 2 + 65536;                    // result u16 + score code u8
constexpr int kLutBytes = 256 * 8;                                // byte (two cells) -> two int32 tile values
constexpr int kScoreLutBytes = 256 * 4;                           // score code -> score
constexpr int kStepSmem = kTableBytes + kStepBlock * 8 + kLutBytes + kScoreLutBytes;  // + one packed board per thread + the LUTs

// The byte LUT's 8-byte entries sit in 16 bank pairs chosen by the low nibble of the index, i.e. by the
// first cell's exponent alone; with small exponents most of a half-warp lands on 3-4 pairs.  Folding the
// second cell's low two bits into bits 2-3 (a bijection on bytes) spreads them over all 16 (7.0 -> 4.1
// wavefronts per read on boards from random play).  The staged board is swizzled once per board instead.
__device__ __forceinline__ uint32_t tfe_lut_slot(uint32_t byte) { return byte ^ ((byte >> 2) & 0x0Cu); }
__device__ __forceinline__ uint64_t tfe_lut_swizzle_board(uint64_t b) {
  uint32_t lo = (uint32_t)b, hi = (uint32_t)(b >> 32);
  lo ^= (lo >> 2) & 0x0C0C0C0Cu;
  hi ^= (hi >> 2) & 0x0C0C0C0Cu;
  return ((uint64_t)hi << 32) | lo;
}

template <int SIDE>
__device__ __forceinline__ void tfe_warp_write_obs(const uint64_t* __restrict__ wboards, const uint2* __restrict__ lut,
                                                   const uint32_t wb, const uint32_t n, int32_t* __restrict__ obs_out) {
  const int lane = threadIdx.x & 31;
  const int rows = (int)((n - wb) < 32u ? (n - wb) : 32u);
  constexpr int W = SIDE * SIDE;
  if (SIDE == 4) {
    // int4 index q = board * 4 + board row, and the staged boards are little-endian, so 16-bit word q of the
    // staging slot IS that row: one conflict-free LDS.U16, then one LUT read per byte (two cells each) --
    // LSU-pipe work instead of twelve shift/mask ops on the ALU pipe
    const uint16_t* wrows = reinterpret_cast<const uint16_t*>(wboards);
    int4* dst = reinterpret_cast<int4*>(obs_out + (size_t)wb * W) + lane;
    if (rows == 32) {                                               // full warp (all but the last): no bounds checks
#pragma unroll
      for (int it = 0; it < 4; ++it) {
        const uint32_t r16 = wrows[it * 32 + lane];
        const uint2 a = lut[r16 & 0xFFu], b = lut[r16 >> 8];
        dst[it * 32] = make_int4((int)a.x, (int)a.y, (int)b.x, (int)b.y);
      }
    } else {
#pragma unroll
      for (int it = 0; it < 4; ++it) {
        if (it * 32 + lane < rows * 4) {
          const uint32_t r16 = wrows[it * 32 + lane];
          const uint2 a = lut[r16 & 0xFFu], b = lut[r16 >> 8];
          dst[it * 32] = make_int4((int)a.x, (int)a.y, (int)b.x, (int)b.y);
        }
      }
    }
  } else {
    int32_t* dst = obs_out + (size_t)wb * W;
#pragma unroll
    for (int it = 0; it < W; ++it) {
      const int q = it * 32 + lane;
      if (q < rows * W) {
        const int r = q / W, c = q - r * W;
        const uint32_t e = (uint32_t)(wboards[r] >> (16 * (c / SIDE) + 4 * (c % SIDE))) & 15u;
        dst[q] = (int)((1u << e) & ~1u);
      }
    }
  }
}

template <int SIDE, typename ActT>
__global__ void __launch_bounds__(kStepBlock, 1)
tfe_step_kernel(uint64_t* __restrict__ boards, const ActT* __restrict__ actions, const uint2* __restrict__ chance,
                const uint4* __restrict__ reset_chance, const uint4* __restrict__ table, uint64_t seed,
                uint64_t env_base, uint32_t counter, uint32_t* counter_dev, int auto_reset,
                int32_t* __restrict__ obs_out, int32_t* __restrict__ reward_out, uint8_t* done_out,
                int64_t* __restrict__ score, uint32_t n) {
  // board indices are 32-bit (n_envs is an int in the ABI; g + stride stays below 2^32): one IMAD.WIDE per address
  extern __shared__ uint4 smem_raw[];
  // ---- prologue: independent of every earlier kernel on the stream (the table was built and synchronised at
  // init), so under programmatic dependent launch it overlaps the previous kernel's tail ----
  // table -> shared memory by TMA bulk copies (12 x 16 KB), issued by one thread, landing on an mbarrier
  __shared__ uint64_t table_bar;
  if (threadIdx.x == 0) pulse_mbar_init(&table_bar, 1);
  __syncthreads();
  if (threadIdx.x == 0) {
    pulse_mbar_expect_tx(&table_bar, kTableBytes);
    constexpr int kChunk = 16384;
#pragma unroll 1
    for (int off = 0; off < kTableBytes; off += kChunk)
      pulse_bulk_g2s(reinterpret_cast<uint8_t*>(smem_raw) + off, reinterpret_cast<const uint8_t*>(table) + off, kChunk,
                     &table_bar);
  }
  uint64_t* wboards = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(smem_raw) + kTableBytes) + (threadIdx.x & ~31);
  uint2* lut = reinterpret_cast<uint2*>(reinterpret_cast<uint8_t*>(smem_raw) + kTableBytes + kStepBlock * 8);
  uint32_t* score_lut = reinterpret_cast<uint32_t*>(reinterpret_cast<uint8_t*>(smem_raw) + kTableBytes + kStepBlock * 8 + kLutBytes);
  if (threadIdx.x < 256) {                                          // the two small LUTs (published by the barrier below)
    lut[tfe_lut_slot(threadIdx.x)] = make_uint2((1u << (threadIdx.x & 15u)) & ~1u, (1u << (threadIdx.x >> 4)) & ~1u);
    score_lut[threadIdx.x] = tfe_code_score(threadIdx.x);
  }
  // ---- everything below reads or writes what earlier kernels may have written ----
  asm volatile("griddepcontrol.wait;" ::: "memory");
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  const uint32_t dev_ctr = pulse_counter_read(counter_dev);
  counter += dev_ctr;
  const uint32_t stride = gridDim.x * kStepBlock;
  uint32_t g = blockIdx.x * kStepBlock + threadIdx.x;
  // first board's inputs are in flight while the table copy lands
  // done flags are carried as the raw byte and only tested where they are used: testing one right after its load
  // (as `auto_reset && done_out[gn]` did) parks the warp on the DRAM round trip the prefetch exists to hide
  uint64_t b = 0; int64_t total = 0; int a = 0; uint32_t done_raw = 0;
  if (g < n) { b = boards[g]; total = score[g]; a = (int)actions[g]; done_raw = done_out[g]; }
  bool table_ready = false;
  TfeTableRows rows;
  rows.result = reinterpret_cast<const uint16_t*>(smem_raw);
  rows.code = reinterpret_cast<const uint8_t*>(smem_raw) + 65536 * 2;
  rows.score_of_code = score_lut;
  __syncthreads();
  const uint32_t lane = threadIdx.x & 31;
  for (; g - lane < n; g += stride) {                              // warp-uniform trip count
    const bool valid = g < n;
    const uint32_t gn = g + stride;
    uint64_t nb_in = 0; int64_t ntotal = 0; int na = 0; uint32_t ndone_raw = 0;
    if (gn < n) { nb_in = boards[gn]; ntotal = score[gn]; na = (int)actions[gn]; ndone_raw = done_out[gn]; }
    uint64_t nb = 0;
    if (valid) {
      const bool was_done = auto_reset && done_raw != 0;
      // one chance block per board-step: the reset stream for a board being re-dealt, the spawn stream
      // otherwise -- a single Philox call either way (no second instruction stream for finished boards)
      PulseU4 c;
      c.x = c.y = c.z = c.w = 0;
      const bool injected = was_done ? reset_chance != nullptr : chance != nullptr;
      if (injected) {
        if (was_done) { const uint4 v = reset_chance[g]; c.x = v.x; c.y = v.y; c.z = v.z; c.w = v.w; }
        else { const uint2 v = chance[g]; c.x = v.x; c.y = v.y; }
      } else {
        c = pulse_tape(seed, env_base + (uint64_t)g, was_done ? PULSE_STREAM_TFE_RESET : PULSE_STREAM_TFE_SPAWN, counter);
      }
      int32_t reward; bool done;
      const int64_t total0 = total;
      if (!table_ready) { pulse_mbar_wait(&table_bar, 0); table_ready = true; }  // first board only
      nb = tfe_step_fused<SIDE>(b, a, was_done, c, total, reward, done, rows);
      if (nb != b) boards[g] = nb;
      if (total != total0) score[g] = total;
      reward_out[g] = reward;
      done_out[g] = done;
    }
    if (obs_out) {
      wboards[lane] = SIDE == 4 ? tfe_lut_swizzle_board(nb) : nb;
      __syncwarp();
      tfe_warp_write_obs<SIDE>(wboards, lut, g - lane, n, obs_out);
      __syncwarp();
    }
    b = nb_in; total = ntotal; a = na; done_raw = ndone_raw;
  }
  pulse_counter_advance(counter_dev, dev_ctr);
}

// ---- K steps per launch ("rollout") -------------------------------------------------------------------
// For open-loop action tapes (random-policy rollouts, BASELINE config 2) the K actions of a board are known up front:
// the table is staged ONCE per K steps instead of once per step (~3 us of a 22 us step launch), and a board and its
// score stay in registers between its steps (no 32 B of state traffic per step).  A thread plays all K steps of one
// board, then takes its next board; actions are [K, N], outputs step-major [K, N, ...], so the stores of a warp are
// contiguous for every k.  Chance comes from the tape at counter + k, exactly as K single-step launches would draw it.
template <int SIDE, typename ActT>
__global__ void __launch_bounds__(kStepBlock, 1)
tfe_rollout_kernel(uint64_t* __restrict__ boards, const ActT* __restrict__ actions, int K, const uint4* __restrict__ table,
                   uint64_t seed, uint64_t env_base, uint32_t counter, int auto_reset, int32_t* __restrict__ obs_out,
                   int32_t* __restrict__ reward_out, uint8_t* __restrict__ done_out, uint8_t* __restrict__ done_state,
                   int64_t* __restrict__ score, uint32_t n) {
  extern __shared__ uint4 smem_raw[];
  __shared__ uint64_t table_bar;
  if (threadIdx.x == 0) pulse_mbar_init(&table_bar, 1);
  __syncthreads();
  if (threadIdx.x == 0) {
    pulse_mbar_expect_tx(&table_bar, kTableBytes);
    constexpr int kChunk = 16384;
#pragma unroll 1
    for (int off = 0; off < kTableBytes; off += kChunk)
      pulse_bulk_g2s(reinterpret_cast<uint8_t*>(smem_raw) + off, reinterpret_cast<const uint8_t*>(table) + off, kChunk,
                     &table_bar);
  }
  uint64_t* wboards = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(smem_raw) + kTableBytes) + (threadIdx.x & ~31);
  uint2* lut = reinterpret_cast<uint2*>(reinterpret_cast<uint8_t*>(smem_raw) + kTableBytes + kStepBlock * 8);
  uint32_t* score_lut = reinterpret_cast<uint32_t*>(reinterpret_cast<uint8_t*>(smem_raw) + kTableBytes + kStepBlock * 8 + kLutBytes);
  if (threadIdx.x < 256) {
    lut[tfe_lut_slot(threadIdx.x)] = make_uint2((1u << (threadIdx.x & 15u)) & ~1u, (1u << (threadIdx.x >> 4)) & ~1u);
    score_lut[threadIdx.x] = tfe_code_score(threadIdx.x);
  }
  TfeTableRows rows;
  rows.result = reinterpret_cast<const uint16_t*>(smem_raw);
  rows.code = reinterpret_cast<const uint8_t*>(smem_raw) + 65536 * 2;
  rows.score_of_code = score_lut;
  __syncthreads();
  pulse_mbar_wait(&table_bar, 0);
  const uint32_t stride = gridDim.x * kStepBlock;
  const uint32_t lane = threadIdx.x & 31;
  const size_t W = (size_t)SIDE * SIDE;
  for (uint32_t g = blockIdx.x * kStepBlock + threadIdx.x; g - lane < n; g += stride) {   // warp-uniform trip count
    const bool valid = g < n;
    uint64_t b = 0; int64_t total = 0; bool was_done = false; int a = 0;
    if (valid) { b = boards[g]; total = score[g]; was_done = done_state[g] != 0; a = (int)actions[g]; }
    for (int k = 0; k < K; ++k) {
      int na = 0;
      if (valid && k + 1 < K) na = (int)actions[(size_t)(k + 1) * n + g];                  // one step ahead
      int32_t reward = 0; bool done = false;
      if (valid) {
        const bool reset_now = auto_reset && was_done;
        const PulseU4 c = pulse_tape(seed, env_base + (uint64_t)g, reset_now ? PULSE_STREAM_TFE_RESET : PULSE_STREAM_TFE_SPAWN,
                                     counter + (uint32_t)k);
        b = tfe_step_fused<SIDE>(b, a, reset_now, c, total, reward, done, rows);
        const size_t o = (size_t)k * n + g;
        reward_out[o] = reward;
        done_out[o] = done;
        was_done = done;
      }
      if (obs_out) {
        wboards[lane] = SIDE == 4 ? tfe_lut_swizzle_board(b) : b;
        __syncwarp();
        tfe_warp_write_obs<SIDE>(wboards, lut, g - lane, n, obs_out + (size_t)k * n * W);
        __syncwarp();
      }
      a = na;
    }
    if (valid) { boards[g] = b; score[g] = total; done_state[g] = was_done; }
  }
}

// builds the row-move table: result[row] (uint16) then code[row] (uint8), 196,608 bytes
__global__ void tfe_build_table_kernel(uint16_t* __restrict__ result, uint8_t* __restrict__ code) {
  const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r < 65536u) {
    uint32_t sc = 0;
    result[tfe_row_slot(r)] = (uint16_t)tfe_squash_row(r, sc);
    code[tfe_row_slot(r)] = (uint8_t)tfe_score_code(r);
  }
}

template <int SIDE>
__global__ void __launch_bounds__(kBlock)
tfe_reset_kernel(uint64_t* __restrict__ boards, int64_t* __restrict__ score, uint8_t* __restrict__ done,
                 const uint4* __restrict__ chance, uint64_t seed, uint64_t env_base, uint32_t counter,
                 int32_t* __restrict__ obs_out, long long n) {
  __shared__ int32_t smem[kBlock * (SIDE * SIDE + 1)];
  for (long long base = (long long)blockIdx.x * kBlock; base < n; base += (long long)gridDim.x * kBlock) {
    const long long g = base + threadIdx.x;
    const bool valid = g < n;
    uint64_t b = 0;
    if (valid) {
      PulseU4 c;
      if (chance) { const uint4 v = chance[g]; c.x = v.x; c.y = v.y; c.z = v.z; c.w = v.w; }
      else c = pulse_tape(seed, env_base + (uint64_t)g, PULSE_STREAM_TFE_RESET, counter);
      b = tfe_reset_board<SIDE>(c);
      boards[g] = b; score[g] = 0; done[g] = 0;
    }
    if (obs_out) tfe_write_obs<SIDE>(smem, b, valid, base, n, obs_out);
  }
}

template <int SIDE>
__global__ void __launch_bounds__(kBlock)
tfe_unpack_kernel(const uint64_t* __restrict__ boards, int32_t* __restrict__ values, long long n) {
  __shared__ int32_t smem[kBlock * (SIDE * SIDE + 1)];
  for (long long base = (long long)blockIdx.x * kBlock; base < n; base += (long long)gridDim.x * kBlock) {
    const long long g = base + threadIdx.x;
    tfe_write_obs<SIDE>(smem, g < n ? boards[g] : 0, g < n, base, n, values);
  }
}

template <int SIDE>
__global__ void __launch_bounds__(kBlock)
tfe_pack_kernel(const int32_t* __restrict__ values, uint64_t* __restrict__ boards, uint8_t* __restrict__ done,
                int* __restrict__ bad, long long n) {
  for (long long g = (long long)blockIdx.x * kBlock + threadIdx.x; g < n; g += (long long)gridDim.x * kBlock) {
    uint64_t b;
    if (!tfe_pack_board<SIDE>(values + g * SIDE * SIDE, b) && bad) atomicAdd(bad, 1);
    boards[g] = b;
    if (done) done[g] = tfe_game_over<SIDE>(b);
  }
}
// Domain audit: the packed board holds tile exponents up to 15 (32768); two 32768 tiles meeting stay unmerged where
// the reference would make 65536 (tfe_body.cuh).  counts[0] += boards holding a 32768 tile, counts[1] += boards
// holding two or more -- the latter are one move away from leaving the packed domain.
__global__ void __launch_bounds__(kBlock)
tfe_domain_kernel(const uint64_t* __restrict__ boards, int* __restrict__ counts, long long n) {
  int one = 0, two = 0;
  for (long long g = (long long)blockIdx.x * kBlock + threadIdx.x; g < n; g += (long long)gridDim.x * kBlock) {
    uint64_t x = boards[g];
    x &= x >> 1; x &= x >> 2; x &= 0x1111111111111111ULL;      // 1 in every nibble that is 15
    const int k = __popcll(x);
    one += k >= 1; two += k >= 2;
  }
  one = __reduce_add_sync(0xFFFFFFFFu, one); two = __reduce_add_sync(0xFFFFFFFFu, two);
  if ((threadIdx.x & 31) == 0) { if (one) atomicAdd(counts, one); if (two) atomicAdd(counts + 1, two); }
}
}  // namespace

#define TFE_SIDE_DISPATCH(side, CALL)                 \
  do {                                                \
    if ((side) == 4) { constexpr int SD = 4; CALL; }  \
    else if ((side) == 3) { constexpr int SD = 3; CALL; } \
    else { constexpr int SD = 2; CALL; }              \
  } while (0)

// ---- per-device row-move table ------------------------------------------------------------------------
#include <atomic>
#include <mutex>
namespace {
std::mutex g_table_mutex;
std::atomic<uint4*> g_table[64];   // published with release / read with acquire: the fast path takes no lock

template <int SIDE, typename ActT>
cudaError_t allow_smem() {
  cudaFuncSetAttribute(tfe_rollout_kernel<SIDE, ActT>, cudaFuncAttributeMaxDynamicSharedMemorySize, kStepSmem);
  return cudaFuncSetAttribute(tfe_step_kernel<SIDE, ActT>, cudaFuncAttributeMaxDynamicSharedMemorySize, kStepSmem);
}

// Builds the table on the current device the first time it is needed (one cudaMalloc, one small kernel,
// one stream synchronisation -- the only call in the library that may synchronise, and only once).
int tfe_ensure_table(cudaStream_t stream, const uint4** out) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= 64) { pulse_set_error("cudaGetDevice failed"); return e ? (int)e : PULSE_EINVAL; }
  uint4* have = g_table[dev].load(std::memory_order_acquire);
  if (!have) {
    // building allocates and synchronises: refuse while the stream is being captured into a CUDA graph
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(stream, &cap) == cudaSuccess && cap != cudaStreamCaptureStatusNone) {
      pulse_set_error("the 2048 row table of this device is not built yet: call pulse_tfe_init before capturing a graph");
      return PULSE_ENOTREADY;
    }
    std::lock_guard<std::mutex> lock(g_table_mutex);
    have = g_table[dev].load(std::memory_order_acquire);
    if (!have) {
      uint4* t = nullptr;
      if ((e = cudaMalloc(&t, kTableBytes)) != cudaSuccess) { pulse_set_error("cudaMalloc(row table): %s", cudaGetErrorString(e)); return (int)e; }
      tfe_build_table_kernel<<<65536 / 256, 256, 0, stream>>>(reinterpret_cast<uint16_t*>(t),
                                                              reinterpret_cast<uint8_t*>(t) + 65536 * 2);
      if ((e = cudaStreamSynchronize(stream)) != cudaSuccess) { pulse_set_error("row table build: %s", cudaGetErrorString(e)); cudaFree(t); return (int)e; }
      allow_smem<4, uint8_t>(); allow_smem<4, int64_t>(); allow_smem<3, uint8_t>(); allow_smem<3, int64_t>();
      allow_smem<2, uint8_t>(); allow_smem<2, int64_t>();
      g_table[dev].store(t, std::memory_order_release);
      have = t;
    }
  }
  *out = have;
  return PULSE_OK;
}
}  // namespace

extern "C" {

int pulse_tfe_init(cudaStream_t stream) {
  const uint4* t;
  return tfe_ensure_table(stream, &t);
}

int pulse_tfe_reset(uint64_t* boards, int64_t* score, uint8_t* done, const uint32_t* chance4_or_null, uint64_t seed,
                    uint64_t env_base, uint32_t counter, int32_t* obs_out_or_null, int n_envs, int side,
                    cudaStream_t stream) {
  PULSE_REQUIRE(n_envs >= 0, "n_envs < 0");
  PULSE_REQUIRE(side >= 2 && side <= 4, "side must be 2, 3 or 4 (square boards; TFE.py:127 only works for squares)");
  if (n_envs == 0) return PULSE_OK;
  PULSE_REQUIRE(boards && score && done, "null buffer");
  const int grid = pulse_grid_for(n_envs, kBlock);
  TFE_SIDE_DISPATCH(side, (tfe_reset_kernel<SD><<<grid, kBlock, 0, stream>>>(
                              boards, score, done, (const uint4*)chance4_or_null, seed, env_base, counter,
                              obs_out_or_null, (long long)n_envs)));
  return pulse_check_launch("pulse_tfe_reset");
}

int pulse_tfe_step(uint64_t* boards, const void* actions, int action_bytes, const uint32_t* chance2_or_null,
                   const uint32_t* reset_chance4_or_null, uint64_t seed, uint64_t env_base, uint32_t counter,
                   uint32_t* counter_dev_or_null, int auto_reset, int32_t* obs_out_or_null, int32_t* reward_out,
                   uint8_t* done_inout, int64_t* score_inout, int n_envs, int side, cudaStream_t stream) {
  PULSE_REQUIRE(n_envs >= 0, "n_envs < 0");
  PULSE_REQUIRE(side >= 2 && side <= 4, "side must be 2, 3 or 4");
  PULSE_REQUIRE(action_bytes == 1 || action_bytes == 8, "actions must be uint8 or int64");
  if (n_envs == 0) return PULSE_OK;
  PULSE_REQUIRE(boards && actions && reward_out && done_inout && score_inout, "null buffer");
  PULSE_REQUIRE(!obs_out_or_null || (reinterpret_cast<uintptr_t>(obs_out_or_null) & 15u) == 0, "obs must be 16-byte aligned");
  const uint4* table = nullptr;
  int rc = tfe_ensure_table(stream, &table);
  if (rc) return rc;
  const long long need = ((long long)n_envs + kStepBlock - 1) / kStepBlock;
  const int grid = (int)(need < pulse_sm_count() ? need : pulse_sm_count());  // one persistent CTA per SM
  // Programmatic dependent launch: the kernel's prologue (table copy into shared memory, LUT) touches nothing an
  // earlier kernel on the stream writes, so it may start while that kernel drains; everything else sits behind
  // griddepcontrol.wait in the kernel.  PULSE_PDL=0 launches plainly.
  static const bool pdl = !(getenv("PULSE_PDL") && atoi(getenv("PULSE_PDL")) == 0);
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cudaLaunchConfig_t lc = {};
  lc.gridDim = dim3((unsigned)grid); lc.blockDim = dim3(kStepBlock); lc.dynamicSmemBytes = kStepSmem; lc.stream = stream;
  lc.attrs = attr; lc.numAttrs = pdl ? 1 : 0;
  cudaError_t le = cudaSuccess;
#define TFE_STEP(ACT)                                                                                          \
  TFE_SIDE_DISPATCH(side, (le = cudaLaunchKernelEx(&lc, tfe_step_kernel<SD, ACT>,                              \
                              boards, (const ACT*)actions, (const uint2*)chance2_or_null,                      \
                              (const uint4*)reset_chance4_or_null, table, seed, env_base, counter,             \
                              counter_dev_or_null, auto_reset, obs_out_or_null, reward_out, done_inout,        \
                              score_inout, (uint32_t)n_envs)))
  if (action_bytes == 1) TFE_STEP(uint8_t);
  else TFE_STEP(int64_t);
#undef TFE_STEP
  if (le != cudaSuccess) { pulse_set_error("pulse_tfe_step launch: %s", cudaGetErrorString(le)); return (int)le; }
  return pulse_check_launch("pulse_tfe_step");
}

int pulse_tfe_unpack(const uint64_t* boards, int32_t* values_out, int n_envs, int side, cudaStream_t stream) {
  PULSE_REQUIRE(n_envs >= 0 && side >= 2 && side <= 4, "bad n_envs / side");
  if (n_envs == 0) return PULSE_OK;
  PULSE_REQUIRE(boards && values_out, "null buffer");
  TFE_SIDE_DISPATCH(side, (tfe_unpack_kernel<SD><<<pulse_grid_for(n_envs, kBlock), kBlock, 0, stream>>>(
                              boards, values_out, (long long)n_envs)));
  return pulse_check_launch("pulse_tfe_unpack");
}

int pulse_tfe_rollout(uint64_t* boards, const void* actions, int action_bytes, int n_steps, uint64_t seed,
                      uint64_t env_base, uint32_t counter, int auto_reset, int32_t* obs_out_or_null, int32_t* reward_out,
                      uint8_t* done_out, uint8_t* done_inout, int64_t* score_inout, int n_envs, int side,
                      cudaStream_t stream) {
  PULSE_REQUIRE(n_envs >= 0 && n_steps >= 0, "n_envs / n_steps < 0");
  PULSE_REQUIRE(side >= 2 && side <= 4, "side must be 2, 3 or 4");
  PULSE_REQUIRE(action_bytes == 1 || action_bytes == 8, "actions must be uint8 or int64");
  if (n_envs == 0 || n_steps == 0) return PULSE_OK;
  PULSE_REQUIRE(boards && actions && reward_out && done_out && done_inout && score_inout, "null buffer");
  PULSE_REQUIRE(!obs_out_or_null || (reinterpret_cast<uintptr_t>(obs_out_or_null) & 15u) == 0, "obs must be 16-byte aligned");
  const uint4* table = nullptr;
  int rc = tfe_ensure_table(stream, &table);
  if (rc) return rc;
  const long long need = ((long long)n_envs + kStepBlock - 1) / kStepBlock;
  const int grid = (int)(need < pulse_sm_count() ? need : pulse_sm_count());
#define TFE_ROLL(ACT)                                                                                                  \
  TFE_SIDE_DISPATCH(side, (tfe_rollout_kernel<SD, ACT><<<grid, kStepBlock, kStepSmem, stream>>>(                        \
                              boards, (const ACT*)actions, n_steps, table, seed, env_base, counter, auto_reset,         \
                              obs_out_or_null, reward_out, done_out, done_inout, score_inout, (uint32_t)n_envs)))
  if (action_bytes == 1) TFE_ROLL(uint8_t); else TFE_ROLL(int64_t);
#undef TFE_ROLL
  return pulse_check_launch("pulse_tfe_rollout");
}

int pulse_tfe_domain_check(const uint64_t* boards, int* counts2, int n_envs, cudaStream_t stream) {
  PULSE_REQUIRE(n_envs >= 0, "n_envs < 0");
  if (n_envs == 0) return PULSE_OK;
  PULSE_REQUIRE(boards && counts2, "null buffer");
  tfe_domain_kernel<<<pulse_grid_for(n_envs, kBlock), kBlock, 0, stream>>>(boards, counts2, (long long)n_envs);
  return pulse_check_launch("pulse_tfe_domain_check");
}

int pulse_tfe_pack(const int32_t* values, uint64_t* boards_out, uint8_t* done_out_or_null, int* bad_count_or_null,
                   int n_envs, int side, cudaStream_t stream) {
  PULSE_REQUIRE(n_envs >= 0 && side >= 2 && side <= 4, "bad n_envs / side");
  if (n_envs == 0) return PULSE_OK;
  PULSE_REQUIRE(values && boards_out, "null buffer");
  TFE_SIDE_DISPATCH(side, (tfe_pack_kernel<SD><<<pulse_grid_for(n_envs, kBlock), kBlock, 0, stream>>>(
                              values, boards_out, done_out_or_null, bad_count_or_null, (long long)n_envs)));
  return pulse_check_launch("pulse_tfe_pack");
}
}
