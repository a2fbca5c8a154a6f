// Tetris kernels and their C ABI (include/pulse_envs.h).  Rules: tetris_body.cuh / DESIGN.md section 6.
// State is rows uint16 [N][20] row-major (40 B per env; this tensor IS the observation, as the
// reference's TFE returns its board array itself), piece uint8 [N], lines int32 [N], done uint8 [N].
// One thread per env; each block copies its contiguous [kBlock][20] slab of rows into shared memory
// with 16 B coalesced loads, threads drop / lock / clear on their own 20 rows there (dynamic row
// indexing stays out of local memory), and the slab goes back with 16 B coalesced stores.  The 28-entry
// shape table is staged in shared memory too.  HBM-bound: DESIGN.md 4.4.
#include "tetris_body.cuh"

namespace {
constexpr int kBlock = 128;
constexpr int kSlabWords = kBlock * TET_H / 2;  // uint32 words in a full slab

__constant__ uint32_t c_shapes[TET_SHAPES] = TET_SHAPE_TABLE;

__device__ __forceinline__ void slab_copy(uint32_t* dst, const uint32_t* src, int words) {
  // full slabs: 16 B per lane; the last slab (any size) falls back to 4 B
  if (words == kSlabWords) {
    const uint4* s4 = reinterpret_cast<const uint4*>(src);
    uint4* d4 = reinterpret_cast<uint4*>(dst);
#pragma unroll
    for (int i = threadIdx.x; i < kSlabWords / 4; i += kBlock) d4[i] = s4[i];
  } else {
    for (int i = threadIdx.x; i < words; i += kBlock) dst[i] = src[i];
  }
}

template <typename ActT>
__global__ void __launch_bounds__(kBlock)
tetris_step_kernel(uint16_t* __restrict__ rows, uint8_t* __restrict__ piece, int32_t* __restrict__ lines,
                   uint8_t* __restrict__ done, const ActT* __restrict__ actions, const uint8_t* __restrict__ next_piece,
                   const uint8_t* __restrict__ reset_piece, uint64_t seed, uint64_t env_base, uint32_t counter,
                   uint32_t* counter_dev, int auto_reset, int32_t* __restrict__ reward_out, long long n) {
  __shared__ __align__(16) uint32_t slab[kSlabWords];
  __shared__ uint32_t shapes[TET_SHAPES];
  const uint32_t dev_ctr = pulse_counter_read(counter_dev);
  counter += dev_ctr;
  if (threadIdx.x < TET_SHAPES) shapes[threadIdx.x] = c_shapes[threadIdx.x];
  for (long long base = (long long)blockIdx.x * kBlock; base < n; base += (long long)gridDim.x * kBlock) {
    const int cnt = (int)((n - base) < kBlock ? (n - base) : kBlock);
    uint32_t* gslab = reinterpret_cast<uint32_t*>(rows + base * TET_H);
    slab_copy(slab, gslab, cnt * TET_H / 2);
    const long long g = base + threadIdx.x;
    // per-env scalars are fetched while the slab copy is in flight
    int pc = 0, nx = 0, rp = 0; int32_t ln = 0; bool dn = false; long long a = 0;
    if (threadIdx.x < cnt) {
      pc = piece[g]; ln = lines[g]; dn = done[g]; a = (long long)actions[g];
      if (next_piece && reset_piece) { nx = next_piece[g]; rp = reset_piece[g]; }
      else {
        const PulseU4 blk = pulse_tape(seed, env_base + (uint64_t)g, PULSE_STREAM_TETRIS_PIECE, counter);
        nx = next_piece ? next_piece[g] : (int)pulse_mulhi32(blk.x, 7u);
        rp = reset_piece ? reset_piece[g] : (int)pulse_mulhi32(blk.y, 7u);
      }
    }
    __syncthreads();
    if (threadIdx.x < cnt) {
      const int pc0 = pc; const int32_t ln0 = ln; const bool dn0 = dn;
      int32_t reward;
      tet_step_one(reinterpret_cast<uint16_t*>(slab) + threadIdx.x * TET_H, shapes, pc, ln, dn, a, nx, rp, auto_reset,
                   reward);
      if (pc != pc0) piece[g] = (uint8_t)pc;
      if (ln != ln0) lines[g] = ln;
      if (dn != dn0) done[g] = dn;
      reward_out[g] = reward;
    }
    __syncthreads();
    slab_copy(gslab, slab, cnt * TET_H / 2);
    __syncthreads();
  }
  pulse_counter_advance(counter_dev, dev_ctr);
}

__global__ void __launch_bounds__(256)
tetris_reset_kernel(uint16_t* __restrict__ rows, uint8_t* __restrict__ piece, int32_t* __restrict__ lines,
                    uint8_t* __restrict__ done, const uint8_t* __restrict__ first_piece, uint64_t seed,
                    uint64_t env_base, uint32_t counter, long long n) {
  uint32_t* w = reinterpret_cast<uint32_t*>(rows);
  const long long words = n * (TET_H / 2), stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < words; i += stride) w[i] = 0;
  for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < n; g += stride) {
    piece[g] = first_piece ? first_piece[g]
                           : (uint8_t)pulse_mulhi32(pulse_tape(seed, env_base + (uint64_t)g, PULSE_STREAM_TETRIS_PIECE, counter).y, 7u);
    lines[g] = 0; done[g] = 0;
  }
}
}  // namespace

extern "C" {

int pulse_tetris_reset(uint16_t* rows, uint8_t* piece, int32_t* lines, uint8_t* done, const uint8_t* first_piece_or_null,
                       uint64_t seed, uint64_t env_base, uint32_t counter, int n_envs, cudaStream_t stream) {
  PULSE_REQUIRE(n_envs >= 0, "n_envs < 0");
  if (n_envs == 0) return PULSE_OK;
  PULSE_REQUIRE(rows && piece && lines && done, "null buffer");
  tetris_reset_kernel<<<pulse_grid_for((long long)n_envs * 10, 256), 256, 0, stream>>>(
      rows, piece, lines, done, first_piece_or_null, seed, env_base, counter, (long long)n_envs);
  return pulse_check_launch("pulse_tetris_reset");
}

int pulse_tetris_step(uint16_t* rows, uint8_t* piece, int32_t* lines, uint8_t* done, const void* actions,
                      int action_bytes, const uint8_t* next_piece_or_null, const uint8_t* reset_piece_or_null,
                      uint64_t seed, uint64_t env_base, uint32_t counter, uint32_t* counter_dev_or_null,
                      int auto_reset, int32_t* reward_out, int n_envs, cudaStream_t stream) {
  PULSE_REQUIRE(n_envs >= 0, "n_envs < 0");
  PULSE_REQUIRE(action_bytes == 1 || action_bytes == 8, "actions must be uint8 or int64");
  if (n_envs == 0) return PULSE_OK;
  PULSE_REQUIRE(rows && piece && lines && done && actions && reward_out, "null buffer");
  PULSE_REQUIRE((reinterpret_cast<uintptr_t>(rows) & 15u) == 0, "rows must be 16-byte aligned");
  const int grid = pulse_grid_for(n_envs, kBlock);
  if (action_bytes == 1)
    tetris_step_kernel<uint8_t><<<grid, kBlock, 0, stream>>>(rows, piece, lines, done, (const uint8_t*)actions,
                                                             next_piece_or_null, reset_piece_or_null, seed, env_base,
                                                             counter, counter_dev_or_null, auto_reset, reward_out,
                                                             (long long)n_envs);
  else
    tetris_step_kernel<int64_t><<<grid, kBlock, 0, stream>>>(rows, piece, lines, done, (const int64_t*)actions,
                                                             next_piece_or_null, reset_piece_or_null, seed, env_base,
                                                             counter, counter_dev_or_null, auto_reset, reward_out,
                                                             (long long)n_envs);
  return pulse_check_launch("pulse_tetris_step");
}
}
