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
  const int grid = (int)(need < pulse_sm_count() * 4 ? need : pulse_sm_count() * 4);
  if (reward_is_float) stats_kernel<float><<<grid, kBlock, 0, stream>>>((const float*)reward, done_or_null, stats4, n);
  else stats_kernel<int32_t><<<grid, kBlock, 0, stream>>>((const int32_t*)reward, done_or_null, stats4, n);
  return pulse_check_launch("pulse_stats_accumulate");
}

// ---- hand-level statistics ------------------------------------------------------
// What the reference's performance loop gathers on the host every step with boolean-mask indexing
// (/root/reference/scripts/Poker/trainGPU_performance.py:188-206: newly_done = dones & ~terminated_before,
// final - initial stack of the Q seat, terminal stage, button-relative position) and reduces afterwards
// (/root/reference/utils/performance.py:104-167: mean delta, win rate, hands, BB/100, LCB95; :170-196 stage
// buckets and street win percentages; :55-59, :199-253 positions, position win rates, seat-balanced BB/100).
// Here one pass per step folds the newly finished hands into a 67-double vector on the device -- the vector
// that is summed over ranks by NCCL:
//   [0] hands  [1] sum d  [2] sum d^2  [3] #(d > 0)
//   per terminal-stage bucket b (0 preflop .. 3 river, 4 = stage >= 4): [4+b] hands  [9+b] sum d  [14+b] wins
//   per position p = (q_seat - button) mod active_players:              [19+p] hands [35+p] sum d [51+p] wins
// Deltas are integers (chips), so every sum is exact in double and independent of the order of the atomics.
namespace {
constexpr int kHandStats = 67;

struct CompatHandView {   // reference-shaped tensors
  const int32_t* stacks; int P; const int32_t* stages; const int32_t* button;
  __device__ int q_stack(long long g, int q) const { return stacks[g * P + q]; }
  __device__ int stage(long long g) const { return stages[g]; }
  __device__ int btn(long long g) const { return button[g]; }
};
struct PackedHandView {   // packed layout: seat-major stacks, stage/button in the table word
  const int32_t* stack; long long n; const uint32_t* tbl;
  __device__ int q_stack(long long g, int q) const { return stack[(long long)q * n + g]; }
  __device__ int stage(long long g) const { return (int)(tbl[4 * g] & 7u); }
  __device__ int btn(long long g) const { return (int)((tbl[4 * g] >> 11) & 15u); }
};

// Accumulation is in 64-bit integers in shared memory (native atomics; a double atomicAdd in shared memory is a
// compare-and-swap loop and every finishing table of a block hits the same few slots), converted to double for
// the one global atomic per non-zero slot per block.  d^2 summed over a 4M-table batch stays far below 2^63.
template <class View>
__global__ void __launch_bounds__(kBlock)
hand_stats_kernel(View v, const uint8_t* __restrict__ done, uint8_t* __restrict__ terminated,
                  int32_t* __restrict__ initial_q, int q_seat, int active_players, int rearm_starting, int rearm_max,
                  double* __restrict__ stats, long long n) {
  // totals live in registers and are reduced by shuffles; the per-bucket / per-position slots are private to each
  // warp, so an atomic only ever meets lanes of its own warp instruction
  __shared__ unsigned long long sh[kBlock / 32][kHandStats];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < (kBlock / 32) * kHandStats; i += kBlock) (&sh[0][0])[i] = 0ull;
  __syncthreads();
  unsigned long long t_n = 0, t_s = 0, t_s2 = 0, t_w = 0;
  unsigned long long* mine = sh[warp];
  for (long long g = (long long)blockIdx.x * kBlock + threadIdx.x; g < n; g += (long long)gridDim.x * kBlock) {
    // rearm (per-table auto-reset): done is raised for one step only, the next deal starts the next hand
    const bool d = done[g] != 0, before = rearm_starting > 0 ? false : terminated[g] != 0;
    if (d && !before) {                                     // newly_done (:190)
      const int fin = v.q_stack(g, q_seat);
      const long long delta = (long long)fin - (long long)initial_q[g];     // :193-194
      int bucket = v.stage(g);
      bucket = bucket >= 4 ? 4 : (bucket < 0 ? 0 : bucket);                 // bucketize_terminal_stages
      int pos = (q_seat - v.btn(g)) % active_players;                        // calculate_q_seat_positions
      if (pos < 0) pos += active_players;
      pos &= 15;
      const unsigned long long ud = (unsigned long long)delta;               // two's complement: sums wrap back
      t_n += 1ull; t_s += ud; t_s2 += (unsigned long long)(delta * delta);
      atomicAdd(&mine[4 + bucket], 1ull); atomicAdd(&mine[9 + bucket], ud);
      atomicAdd(&mine[19 + pos], 1ull); atomicAdd(&mine[35 + pos], ud);
      if (delta > 0) { t_w += 1ull; atomicAdd(&mine[14 + bucket], 1ull); atomicAdd(&mine[51 + pos], 1ull); }
      if (rearm_starting > 0) {   // the next hand's starting stack: what the deal will leave in the seat (PokerGPU.py:105-108)
        initial_q[g] = (fin == 0 || fin > rearm_max) ? rearm_starting : fin;
      } else {
        terminated[g] = 1;                                  // terminated |= dones (:187)
      }
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    t_n += __shfl_down_sync(0xFFFFFFFFu, t_n, o); t_s += __shfl_down_sync(0xFFFFFFFFu, t_s, o);
    t_s2 += __shfl_down_sync(0xFFFFFFFFu, t_s2, o); t_w += __shfl_down_sync(0xFFFFFFFFu, t_w, o);
  }
  if (lane == 0) { mine[0] = t_n; mine[1] = t_s; mine[2] = t_s2; mine[3] = t_w; }
  __syncthreads();
  for (int i = threadIdx.x; i < kHandStats; i += kBlock) {
    unsigned long long acc = 0ull;
#pragma unroll
    for (int w = 0; w < kBlock / 32; ++w) acc += sh[w][i];
    if (acc != 0ull) atomicAdd(stats + i, (double)(long long)acc);
  }
}

template <class View>
int launch_hand_stats(View v, const uint8_t* done, uint8_t* terminated, int32_t* initial_q, int q_seat,
                      int active_players, int rearm_starting, int rearm_max, double* stats, long long n,
                      cudaStream_t stream, const char* what) {
  long long need = (n + kBlock - 1) / kBlock;
  const int grid = (int)(need < pulse_sm_count() * 8 ? need : pulse_sm_count() * 8);
  hand_stats_kernel<View><<<grid, kBlock, 0, stream>>>(v, done, terminated, initial_q, q_seat, active_players,
                                                       rearm_starting, rearm_max, stats, n);
  return pulse_check_launch(what);
}
}  // namespace

extern "C" int pulse_poker_hand_stats(const int32_t* stacks, int n_players, const int32_t* stages,
                                      const int32_t* button, const uint8_t* done, uint8_t* terminated_inout,
                                      int32_t* initial_q_stacks, int q_seat, int active_players,
                                      double* stats67, long long n_games, cudaStream_t stream) {
  PULSE_REQUIRE(n_games >= 0, "n_games < 0");
  PULSE_REQUIRE(n_players >= 2 && n_players <= 16 && active_players >= 1 && active_players <= n_players, "bad table size");
  PULSE_REQUIRE(q_seat >= 0 && q_seat < n_players, "q_seat out of range");
  if (n_games == 0) return PULSE_OK;
  PULSE_REQUIRE(stacks && stages && button && done && terminated_inout && initial_q_stacks && stats67, "null buffer");
  return launch_hand_stats(CompatHandView{stacks, n_players, stages, button}, done, terminated_inout, initial_q_stacks,
                           q_seat, active_players, 0, 0, stats67, n_games, stream, "pulse_poker_hand_stats");
}

extern "C" int pulse_poker_packed_hand_stats(const PulsePokerPacked* st, const PulsePokerConfig* cfg,
                                             const uint8_t* done, uint8_t* terminated_inout,
                                             int32_t* initial_q_stacks, int q_seat, int rearm, double* stats67,
                                             cudaStream_t stream) {
  PULSE_REQUIRE(st && cfg, "null state/config");
  PULSE_REQUIRE(q_seat >= 0 && q_seat < cfg->n_players, "q_seat out of range");
  PULSE_REQUIRE(cfg->active_players >= 1 && cfg->active_players <= cfg->n_players, "bad active_players");
  if (cfg->n_games == 0) return PULSE_OK;
  PULSE_REQUIRE(done && terminated_inout && initial_q_stacks && stats67, "null buffer");
  return launch_hand_stats(PackedHandView{st->stack, (long long)cfg->n_games, st->tbl}, done, terminated_inout,
                           initial_q_stacks, q_seat, cfg->active_players, rearm ? cfg->starting_bbs : 0, cfg->max_bbs,
                           stats67, cfg->n_games, stream, "pulse_poker_packed_hand_stats");
}
