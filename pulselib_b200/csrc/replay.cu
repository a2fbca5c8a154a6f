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

#include <atomic>

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

// Row mover.  VEC: obs_width % 4 == 0 and 16 B aligned buffers -> a row is L4 = W/4 float4 chunks and a warp moves
// 32 / L4 selected tables per pass (W = 16: eight tables per pass, every lane busy); otherwise one table per pass,
// 4 B per lane.  Slots advance by one per selected table, so the ring position is reduced modulo the capacity once
// per warp and then kept in range with a compare-and-subtract (no 64-bit division per row).
template <bool VEC>
__global__ void __launch_bounds__(kBlock)
replay_write_kernel(const float* __restrict__ state, const float* __restrict__ next_state, int W,
                    const int64_t* __restrict__ actions, const float* __restrict__ rewards,
                    const uint8_t* __restrict__ dones, const int32_t* __restrict__ seat,
                    const uint8_t* __restrict__ term_before, int q_seat, const int32_t* __restrict__ offsets,
                    const unsigned long long* __restrict__ base, float* __restrict__ state_buf,
                    float* __restrict__ next_buf, int64_t* __restrict__ action_buf, float* __restrict__ reward_buf,
                    uint8_t* __restrict__ done_buf, long long capacity, long long n) {
  __shared__ int warp_cnt[kBlock / 32];
  __shared__ int src_list[kBlock / 32][32];
  const long long g = (long long)blockIdx.x * kBlock + threadIdx.x;
  const bool sel = rp_selected(seat, term_before, q_seat, g, n);
  const unsigned bal = __ballot_sync(0xFFFFFFFFu, sel);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int cnt = __popc(bal);
  if (lane == 0) warp_cnt[warp] = cnt;
  if (sel) src_list[warp][__popc(bal & ((1u << lane) - 1u))] = threadIdx.x;   // rank among the warp's selected
  __syncthreads();
  if (cnt == 0) return;
  int woff = 0;
  for (int w = 0; w < warp; ++w) woff += warp_cnt[w];
  const long long first = (long long)((*base + (unsigned long long)(offsets[blockIdx.x] + woff)) %
                                      (unsigned long long)capacity);
  const long long blk = (long long)blockIdx.x * kBlock;
  if (lane < cnt) {                                    // scalars: lane k moves the k-th selected table's
    const long long src = blk + src_list[warp][lane];
    long long slot = first + lane;
    while (slot >= capacity) slot -= capacity;
    action_buf[slot] = actions[src];
    reward_buf[slot] = rewards[src];
    done_buf[slot] = dones[src];
  }
  if (VEC) {
    const int L4 = W >> 2, per_pass = 32 / L4;         // W <= 128 is checked by the launcher
    const int r = lane / L4, c = lane - r * L4;
    for (int k0 = 0; k0 < cnt; k0 += per_pass) {
      const int k = k0 + r;
      if (r < per_pass && k < cnt) {
        const long long src = blk + src_list[warp][k];
        long long slot = first + k;
        while (slot >= capacity) slot -= capacity;
        reinterpret_cast<float4*>(state_buf + slot * W)[c] = reinterpret_cast<const float4*>(state + src * W)[c];
        reinterpret_cast<float4*>(next_buf + slot * W)[c] = reinterpret_cast<const float4*>(next_state + src * W)[c];
      }
    }
  } else {
    for (int k = 0; k < cnt; ++k) {
      const long long src = blk + src_list[warp][k];
      long long slot = first + k;
      while (slot >= capacity) slot -= capacity;
      for (int c = lane; c < W; c += 32) {
        state_buf[slot * W + c] = state[src * W + c];
        next_buf[slot * W + c] = next_state[src * W + c];
      }
    }
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
  const bool vec = (obs_width & 3) == 0 && obs_width <= 128 &&
                   ((reinterpret_cast<uintptr_t>(state) | reinterpret_cast<uintptr_t>(next_state) |
                     reinterpret_cast<uintptr_t>(state_buf) | reinterpret_cast<uintptr_t>(next_state_buf)) & 15u) == 0;
#define RP_WRITE(V)                                                                                              \
  replay_write_kernel<V><<<(int)nb, kBlock, 0, stream>>>(state, next_state, obs_width, actions, rewards, dones,   \
                                                         seat_idx_before, terminated_before_or_null, q_seat, counts, \
                                                         base, state_buf, next_state_buf, action_buf, reward_buf,  \
                                                         done_buf, capacity, n_games)
  if (vec) RP_WRITE(true); else RP_WRITE(false);
#undef RP_WRITE
  return pulse_check_launch("pulse_replay_append");
}

// ---- pulse_replay_append_static: the same append with NO allocation anywhere ------------------------------------
// For callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
// it): the scratch of the three launches above lives in a file-scope device array instead of caller memory.  It has no
// counterpart in the reference.
//   * limit: n_games <= 4,194,304 (the bench's per-GPU batch); the array is 65,792 B and every GPU that loads the
//     library holds it until the process exits;
//   * every call on a device shares the array: calls that overlap on the GPU (different streams, no dependency) race.
//     Nothing in it has to start at zero -- the count kernel writes counts[0 .. nb) and the scan kernel writes the base
//     word before anything reads them -- so the last call's contents are simply overwritten;
//   * in host code `g_replay_static_scratch` is a host-side stand-in, not a device address: the address the kernels
//     get comes from cudaGetSymbolAddress, resolved once per device with that device current.
namespace {
constexpr long long kStaticMaxGames = 4194304;
constexpr size_t kStaticScratchBytes = (size_t)(kStaticMaxGames / kBlock) * 4 + 256;   // counts + the base word, padded
static __device__ __align__(256) unsigned char g_replay_static_scratch[kStaticScratchBytes];
constexpr int kStaticMaxDevices = 64;
std::atomic<void*> g_static_scratch_ptr[kStaticMaxDevices];   // zero-initialised: not resolved yet

int replay_static_scratch(void** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) { pulse_set_error("pulse_replay_append_static: cudaGetDevice: %s", cudaGetErrorString(e)); return (int)e; }
  if (dev < 0 || dev >= kStaticMaxDevices) { pulse_set_error("pulse_replay_append_static: device ordinal %d out of range", dev); return PULSE_EINVAL; }
  void* p = g_static_scratch_ptr[dev].load(std::memory_order_acquire);
  if (p == nullptr) {
    e = cudaGetSymbolAddress(&p, g_replay_static_scratch);
    if (e != cudaSuccess || p == nullptr) {
      pulse_set_error("pulse_replay_append_static: cudaGetSymbolAddress on device %d: %s", dev, cudaGetErrorString(e));
      return e != cudaSuccess ? (int)e : PULSE_EINVAL;
    }
    g_static_scratch_ptr[dev].store(p, std::memory_order_release);   // two threads racing here store the same address
  }
  *out = p;
  return PULSE_OK;
}
}  // namespace

extern "C" long long pulse_replay_static_max_games(void) { return kStaticMaxGames; }

extern "C" int pulse_replay_append_static(const float* state, const float* next_state, int obs_width,
                                          const int64_t* actions, const float* rewards, const uint8_t* dones,
                                          const int32_t* seat_idx_before, const uint8_t* terminated_before_or_null,
                                          int q_seat, float* state_buf, float* next_state_buf, int64_t* action_buf,
                                          float* reward_buf, uint8_t* done_buf, long long capacity,
                                          uint64_t* total_inout, uint64_t* appended_out, long long n_games,
                                          cudaStream_t stream) {
  PULSE_REQUIRE(n_games >= 0, "bad sizes");
  if (n_games > kStaticMaxGames) {
    pulse_set_error("pulse_replay_append_static: n_games = %lld exceeds the static scratch's limit of %lld tables per call; "
                    "use pulse_replay_append with caller scratch", n_games, kStaticMaxGames);
    return PULSE_EINVAL;
  }
  static_assert(kStaticScratchBytes >= (size_t)(kStaticMaxGames / kBlock) * 4 + 32, "scratch holds counts + base");
  void* scratch = nullptr;
  const int rc = replay_static_scratch(&scratch);
  if (rc != PULSE_OK) return rc;
  return pulse_replay_append(state, next_state, obs_width, actions, rewards, dones, seat_idx_before,
                             terminated_before_or_null, q_seat, state_buf, next_state_buf, action_buf, reward_buf,
                             done_buf, capacity, total_inout, appended_out, scratch, n_games, stream);
}
