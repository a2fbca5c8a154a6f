// Transition hand-off to the learner: the rows of this step that belong to the learning
// seat are appended, in table order, to a device ring buffer.
//
// The reference does this on the host side of every step with five boolean-mask gathers and a clone
// (/root/reference/scripts/Poker/trainGPU.py:79-94: q_mask = seat_idx == q_seat, active = q_mask & ~terminated,
// state[active], actions[active], rewards[active], next_state[active], dones[active]) and its replay buffer
// appends a batch at `arange(position, position + batch) % capacity`
// (/root/reference/utils/ReplayBuffer.py:47-60).  Here: count per 256-table block, one-block exclusive scan,
// then a warp-cooperative copy -- every selected table's two observation rows are moved by all 32 lanes
// (coalesced 4 B x 32), its scalars by lane 0.  The total number of transitions ever appended lives on the
// device (uint64), so nothing synchronises with the host; position = total % capacity, size = min(total, capacity).
// Bytes: 2 reads of 1-4 B per table to select, then 2 x (W x 4 B read + write) + 13 B per selected table.
#include "common.cuh"

namespace {
constexpr int kBlock = 256;

__device__ __forceinline__ bool rp_selected(const int32_t* __restrict__ seat, const uint8_t* __restrict__ term_before,
                                            int q_seat, long long g, long long n) {
  return g < n && seat[g] == q_seat && (term_before == nullptr || term_before[g] == 0);
}

__global__ void __launch_bounds__(kBlock)
replay_count_kernel(const int32_t* __restrict__ seat, const uint8_t* __restrict__ term_before, int q_seat,
                    int32_t* __restrict__ counts, long long n) {
  const long long g = (long long)blockIdx.x * kBlock + threadIdx.x;
  const int c = __syncthreads_count(rp_selected(seat, term_before, q_seat, g, n));
  if (threadIdx.x == 0) counts[blockIdx.x] = c;
}

// exclusive scan of counts[0..nb) in place (one block); scratch[nb] = ring position before this append (as two
// words), total advanced.
__global__ void __launch_bounds__(1024)
replay_scan_kernel(int32_t* __restrict__ counts, int nb, unsigned long long* __restrict__ total,
                   unsigned long long* __restrict__ base_out, unsigned long long* __restrict__ appended_out) {
  __shared__ int warp_sum[32];
  __shared__ int carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (int start = 0; start < nb; start += 1024) {
    const int i = start + threadIdx.x;
    const int v = i < nb ? counts[i] : 0;
    int x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int y = __shfl_up_sync(0xFFFFFFFFu, x, o);
      if ((threadIdx.x & 31) >= o) x += y;
    }
    if ((threadIdx.x & 31) == 31) warp_sum[threadIdx.x >> 5] = x;
    __syncthreads();
    if (threadIdx.x < 32) {
      int w = warp_sum[threadIdx.x];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int y = __shfl_up_sync(0xFFFFFFFFu, w, o);
        if (threadIdx.x >= o) w += y;
      }
      warp_sum[threadIdx.x] = w;  // inclusive over warps
    }
    __syncthreads();
    const int before = carry + (threadIdx.x >= 32 ? warp_sum[(threadIdx.x >> 5) - 1] : 0) + x - v;
    if (i < nb) counts[i] = before;
    __syncthreads();
    if (threadIdx.x == 0) carry += warp_sum[31];
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    *base_out = *total;
    *appended_out = (unsigned long long)carry;
    *total += (unsigned long long)carry;
  }
}

__global__ void __launch_bounds__(kBlock)
replay_write_kernel(const float* __restrict__ state, const float* __restrict__ next_state, int W,
                    const int64_t* __restrict__ actions, const float* __restrict__ rewards,
                    const uint8_t* __restrict__ dones, const int32_t* __restrict__ seat,
                    const uint8_t* __restrict__ term_before, int q_seat, const int32_t* __restrict__ offsets,
                    const unsigned long long* __restrict__ base, float* __restrict__ state_buf,
                    float* __restrict__ next_buf, int64_t* __restrict__ action_buf, float* __restrict__ reward_buf,
                    uint8_t* __restrict__ done_buf, long long capacity, long long n) {
  __shared__ int warp_cnt[kBlock / 32];
  const long long g = (long long)blockIdx.x * kBlock + threadIdx.x;
  const bool sel = rp_selected(seat, term_before, q_seat, g, n);
  const unsigned bal = __ballot_sync(0xFFFFFFFFu, sel);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) warp_cnt[warp] = __popc(bal);
  __syncthreads();
  int woff = 0;
  for (int w = 0; w < warp; ++w) woff += warp_cnt[w];
  const unsigned long long first = *base + (unsigned long long)(offsets[blockIdx.x] + woff);
  unsigned m = bal;
  int k = 0;
  while (m) {                                   // all lanes move the k-th selected table of this warp
    const int src_lane = __ffs(m) - 1;
    m &= m - 1;
    const long long src = (long long)blockIdx.x * kBlock + warp * 32 + src_lane;
    const long long slot = (long long)((first + (unsigned long long)k) % (unsigned long long)capacity);
    for (int c = lane; c < W; c += 32) {
      state_buf[slot * W + c] = state[src * W + c];
      next_buf[slot * W + c] = next_state[src * W + c];
    }
    if (lane == 0) {
      action_buf[slot] = actions[src];
      reward_buf[slot] = rewards[src];
      done_buf[slot] = dones[src];
    }
    ++k;
  }
}
}  // namespace

extern "C" size_t pulse_replay_scratch_bytes(long long n_games) {
  const long long nb = (n_games + kBlock - 1) / kBlock;
  return (size_t)(nb * 4 + 32);
}

extern "C" int pulse_replay_append(const float* state, const float* next_state, int obs_width, const int64_t* actions,
                                   const float* rewards, const uint8_t* dones, const int32_t* seat_idx_before,
                                   const uint8_t* terminated_before_or_null, int q_seat, float* state_buf,
                                   float* next_state_buf, int64_t* action_buf, float* reward_buf, uint8_t* done_buf,
                                   long long capacity, uint64_t* total_inout, uint64_t* appended_out, void* scratch,
                                   long long n_games, cudaStream_t stream) {
  PULSE_REQUIRE(n_games >= 0 && capacity > 0 && obs_width > 0, "bad sizes");
  if (n_games == 0) return PULSE_OK;
  PULSE_REQUIRE(state && next_state && actions && rewards && dones && seat_idx_before, "null step tensor");
  PULSE_REQUIRE(state_buf && next_state_buf && action_buf && reward_buf && done_buf && total_inout && appended_out && scratch,
                "null buffer");
  PULSE_REQUIRE(n_games <= capacity, "one step's batch must fit the ring (n_games <= capacity)");
  const long long nb = (n_games + kBlock - 1) / kBlock;
  int32_t* counts = reinterpret_cast<int32_t*>(scratch);
  unsigned long long* base = reinterpret_cast<unsigned long long*>(reinterpret_cast<char*>(scratch) + ((nb * 4 + 15) / 16) * 16);
  replay_count_kernel<<<(int)nb, kBlock, 0, stream>>>(seat_idx_before, terminated_before_or_null, q_seat, counts, n_games);
  replay_scan_kernel<<<1, 1024, 0, stream>>>(counts, (int)nb, reinterpret_cast<unsigned long long*>(total_inout), base,
                                             reinterpret_cast<unsigned long long*>(appended_out));
  replay_write_kernel<<<(int)nb, kBlock, 0, stream>>>(state, next_state, obs_width, actions, rewards, dones,
                                                      seat_idx_before, terminated_before_or_null, q_seat, counts, base,
                                                      state_buf, next_state_buf, action_buf, reward_buf, done_buf,
                                                      capacity, n_games);
  return pulse_check_launch("pulse_replay_append");
}
