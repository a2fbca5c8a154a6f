// 2048 / TFE kernels and their C ABI (include/pulse_envs.h).  Board rules and the packed layout are in
// tfe_body.cuh.  One thread per board, one launch per step: 8 B board in, 8 B out, optional int32
// side x side observation staged through shared memory so the [N, side*side] slab is written with
// coalesced stores.  HBM-bound (106 B per board-step with the int32 observation, 42 B
// without).
#include "tfe_body.cuh"

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

template <int SIDE, typename ActT>
__global__ void __launch_bounds__(kBlock)
tfe_step_kernel(uint64_t* __restrict__ boards, const ActT* __restrict__ actions, const uint2* __restrict__ chance,
                const uint4* __restrict__ reset_chance, uint64_t seed, uint64_t env_base, uint32_t counter,
                uint32_t* counter_dev, int auto_reset, int32_t* __restrict__ obs_out,
                int32_t* __restrict__ reward_out, uint8_t* __restrict__ done_out, int64_t* __restrict__ score,
                long long n) {
  __shared__ int32_t smem[kBlock * (SIDE * SIDE + 1)];
  const uint32_t dev_ctr = pulse_counter_read(counter_dev);
  counter += dev_ctr;
  for (long long base = (long long)blockIdx.x * kBlock; base < n; base += (long long)gridDim.x * kBlock) {
    const long long g = base + threadIdx.x;
    const bool valid = g < n;
    uint64_t b = 0;
    if (valid) {
      b = boards[g];
      const int a = (int)actions[g];
      int64_t total = score[g];
      const bool was_done = auto_reset && tfe_game_over<SIDE>(b);  // done == game_over(board), always
      PulseU4 sp, rc;
      sp.x = sp.y = sp.z = sp.w = 0; rc = sp;
      if (was_done) {
        if (reset_chance) { const uint4 c = reset_chance[g]; rc.x = c.x; rc.y = c.y; rc.z = c.z; rc.w = c.w; }
        else rc = pulse_tape(seed, env_base + (uint64_t)g, PULSE_STREAM_TFE_RESET, counter);
      } else {
        if (chance) { const uint2 c = chance[g]; sp.x = c.x; sp.y = c.y; }
        else sp = pulse_tape(seed, env_base + (uint64_t)g, PULSE_STREAM_TFE_SPAWN, counter);
      }
      int32_t reward; bool done;
      const int64_t total0 = total;
      const uint64_t nb = tfe_step_one<SIDE>(b, a, was_done, auto_reset, sp, rc, total, reward, done);
      if (nb != b) boards[g] = nb;
      if (total != total0) score[g] = total;
      reward_out[g] = reward;
      done_out[g] = done;
      b = nb;
    }
    if (obs_out) tfe_write_obs<SIDE>(smem, b, valid, base, n, obs_out);
  }
  pulse_counter_advance(counter_dev, dev_ctr);
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
tfe_pack_kernel(const int32_t* __restrict__ values, uint64_t* __restrict__ boards, int* __restrict__ bad, long long n) {
  for (long long g = (long long)blockIdx.x * kBlock + threadIdx.x; g < n; g += (long long)gridDim.x * kBlock) {
    uint64_t b;
    if (!tfe_pack_board<SIDE>(values + g * SIDE * SIDE, b) && bad) atomicAdd(bad, 1);
    boards[g] = b;
  }
}
}  // namespace

#define TFE_SIDE_DISPATCH(side, CALL)                 \
  do {                                                \
    if ((side) == 4) { constexpr int SD = 4; CALL; }  \
    else if ((side) == 3) { constexpr int SD = 3; CALL; } \
    else { constexpr int SD = 2; CALL; }              \
  } while (0)

extern "C" {

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
                   uint32_t* counter_dev_or_null, int auto_reset, int32_t* obs_out_or_null, int32_t* reward_out, uint8_t* done_out,
                   int64_t* score_inout, int n_envs, int side, cudaStream_t stream) {
  PULSE_REQUIRE(n_envs >= 0, "n_envs < 0");
  PULSE_REQUIRE(side >= 2 && side <= 4, "side must be 2, 3 or 4");
  PULSE_REQUIRE(action_bytes == 1 || action_bytes == 8, "actions must be uint8 or int64");
  if (n_envs == 0) return PULSE_OK;
  PULSE_REQUIRE(boards && actions && reward_out && done_out && score_inout, "null buffer");
  const int grid = pulse_grid_for(n_envs, kBlock);
#define TFE_STEP(ACT)                                                                                          \
  TFE_SIDE_DISPATCH(side, (tfe_step_kernel<SD, ACT><<<grid, kBlock, 0, stream>>>(                              \
                              boards, (const ACT*)actions, (const uint2*)chance2_or_null,                      \
                              (const uint4*)reset_chance4_or_null, seed, env_base, counter,                    \
                              counter_dev_or_null, auto_reset,                                                 \
                              obs_out_or_null, reward_out, done_out, score_inout, (long long)n_envs)))
  if (action_bytes == 1) TFE_STEP(uint8_t);
  else TFE_STEP(int64_t);
#undef TFE_STEP
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

int pulse_tfe_pack(const int32_t* values, uint64_t* boards_out, int* bad_count_or_null, int n_envs, int side,
                   cudaStream_t stream) {
  PULSE_REQUIRE(n_envs >= 0 && side >= 2 && side <= 4, "bad n_envs / side");
  if (n_envs == 0) return PULSE_OK;
  PULSE_REQUIRE(values && boards_out, "null buffer");
  TFE_SIDE_DISPATCH(side, (tfe_pack_kernel<SD><<<pulse_grid_for(n_envs, kBlock), kBlock, 0, stream>>>(
                              values, boards_out, bad_count_or_null, (long long)n_envs)));
  return pulse_check_launch("pulse_tfe_pack");
}
}
