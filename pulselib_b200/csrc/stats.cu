// Run statistics of a step's outputs, accumulated on the device: the vector
// that is all-reduced over NCCL once per reporting interval.  The reference keeps these on the host with
// a sync per step (scripts/Poker/trainGPU.py:96-108, scripts/Poker/trainGPU_performance.py:188-206).
//   stats[0] += n            env-steps taken
//   stats[1] += #done        episodes (hands / games) that ended on this step, or envs currently done
//   stats[2] += sum(reward)
//   stats[3] += sum(reward^2)
// One pass over reward (4 B) and done (1 B) per env; warp-shuffle reduction, one atomicAdd(double) x3
// per block.
#include "common.cuh"

namespace {
constexpr int kBlock = 256;

template <typename RewT>
__global__ void __launch_bounds__(kBlock)
stats_kernel(const RewT* __restrict__ reward, const uint8_t* __restrict__ done, double* __restrict__ stats, long long n) {
  double s = 0.0, s2 = 0.0;
  unsigned d = 0;
  for (long long i = (long long)blockIdx.x * kBlock + threadIdx.x; i < n; i += (long long)gridDim.x * kBlock) {
    const double r = (double)reward[i];
    s += r; s2 += r * r;
    if (done) d += done[i] != 0;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    s += __shfl_down_sync(0xFFFFFFFFu, s, o);
    s2 += __shfl_down_sync(0xFFFFFFFFu, s2, o);
    d += __shfl_down_sync(0xFFFFFFFFu, d, o);
  }
  __shared__ double sh[3][kBlock / 32];
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  if (l == 0) { sh[0][w] = s; sh[1][w] = s2; sh[2][w] = (double)d; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double a = 0, b = 0, c = 0;
    for (int k = 0; k < kBlock / 32; ++k) { a += sh[0][k]; b += sh[1][k]; c += sh[2][k]; }
    atomicAdd(stats + 1, c); atomicAdd(stats + 2, a); atomicAdd(stats + 3, b);
    if (blockIdx.x == 0) atomicAdd(stats + 0, (double)n);
  }
}
}  // namespace

extern "C" int pulse_stats_accumulate(const void* reward, int reward_is_float, const uint8_t* done_or_null,
                                      double* stats4, long long n, cudaStream_t stream) {
  PULSE_REQUIRE(n >= 0, "n < 0");
  if (n == 0) return PULSE_OK;
  PULSE_REQUIRE(reward && stats4, "null buffer");
  long long need = (n + kBlock - 1) / kBlock;
  const int grid = (int)(need < 148 * 4 ? need : 148 * 4);
  if (reward_is_float) stats_kernel<float><<<grid, kBlock, 0, stream>>>((const float*)reward, done_or_null, stats4, n);
  else stats_kernel<int32_t><<<grid, kBlock, 0, stream>>>((const int32_t*)reward, done_or_null, stats4, n);
  return pulse_check_launch("pulse_stats_accumulate");
}
