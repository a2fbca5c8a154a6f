// PokerGPU with HOST buffers: the end-to-end path (include/pulse_envs.h, "host-buffer PokerGPU").
//
// The reference's step contract -- int64 actions in, float32 [N, 13 + 3 (M - 1)] observations, float32 rewards and
// bool done flags out (/root/reference/environments/Poker/PokerGPU.py:527-633) -- for callers whose arrays live in host
// memory.  What crosses PCIe is not that contract (8 B in, 69 B out per table-step heads-up) but:
//   in   1 B per table (the action id, 0..12) + 4 B per 32 tables (where the warp's records go)
//   out  one packed record (poker_wire.cuh: 28 B heads-up) per table that was LIVE when the step began, compacted
// and a pool of host threads expands the records into the caller's arrays with streaming stores.  A table that was
// already finished repeats its row (the caller's buffer keeps it), reward 0, done 1 -- exactly what the reference
// returns for it.  The batch is cut into chunks, each on its own stream: chunk c+1's copies and step kernel run while
// the threads expand chunk c.
//
// The handle owns its device state (the packed layout of poker_packed_body.cuh), pinned staging buffers, streams and
// threads; the caller's arrays may be pageable.  One handle = one env on one GPU; calls on one handle are serial.
#include "poker_packed_body.cuh"
#include "host_expand.h"
#include "host_pool.h"

#include <emmintrin.h>

#include <atomic>
#include <chrono>
#include <condition_variable>
#include <functional>
#include <mutex>
#include <new>
#include <thread>
#include <vector>

namespace {

constexpr int kBlock = 128;

// ---- device side ------------------------------------------------------------------------------------------------
// Step in wire mode: uint8 actions, no float32 observation / reward / done arrays at all; a table that was live when
// the step began writes its record to slot warp_off[warp] + (rank among the warp's live lanes).  The host knows which
// tables are live (it expanded the previous step's done flags) and sends the offsets with the actions, so the device
// needs no scan, the records are in table order, and the host knows the copy size before the kernel has run.
template <int MAXP, int MT, int MINB>
__global__ void __launch_bounds__(kBlock, MINB)
poker_host_step_kernel(PulsePokerPacked S, PokerParams pr, const uint8_t* __restrict__ actions,
                       const uint32_t* __restrict__ warp_off, uint32_t* __restrict__ wire, int starting_bbs,
                       int max_bbs, long long first, long long last, long long n, uint32_t* __restrict__ epoch_word,
                       uint32_t epoch) {
  if (blockIdx.x == 0 && threadIdx.x == 0) *epoch_word = epoch;   // copied to the host behind the records: "chunk landed"
  // one chunk of the batch: tables first .. last-1 (first a multiple of 1024); the seat-major arrays keep the whole
  // batch's stride n; warp_off and wire are the chunk's own
  constexpr int R = pkw_row_words(MT);
  pr.M = MT;
  const unsigned lane = threadIdx.x & 31u;
  for (long long base = first + (long long)blockIdx.x * kBlock; base < last; base += (long long)gridDim.x * kBlock) {
    const long long g = base + threadIdx.x;
    uint32_t w[R];
    bool live = false;
    uint32_t off = 0;
    if (g < last) {
      off = warp_off[(g - first) >> 5];
      PokerTable<MAXP> t;
      const PackedStepResult r = packed_step_core<MAXP, false, 0, uint8_t>(S, pr, actions, 0, 0, 0, starting_bbs, max_bbs,
                                                                           g, n, nullptr, t);
      live = !r.was_done;
      if (live) pkw_pack_row<MAXP, MT>(t, pr, r.reward, w);
    }
    const unsigned bal = __ballot_sync(0xFFFFFFFFu, live);
    if (live) {
      uint32_t* dst = wire + ((size_t)off + __popc(bal & ((1u << lane) - 1u))) * R;
#pragma unroll
      for (int i = 0; i < R; ++i) dst[i] = w[i];
    }
  }
}

// After a whole-batch deal: every table's record, uncompacted (slot = table), reward 0.
template <int MAXP, int MT>
__global__ void __launch_bounds__(kBlock)
poker_host_rows_kernel(PulsePokerPacked S, PokerParams pr, uint32_t* __restrict__ wire, long long n,
                       uint32_t* __restrict__ epoch_words, int n_words, uint32_t epoch) {
  if (blockIdx.x == 0 && (int)threadIdx.x < n_words) epoch_words[threadIdx.x] = epoch;
  constexpr int R = pkw_row_words(MT);
  pr.M = MT;
  for (long long g = (long long)blockIdx.x * kBlock + threadIdx.x; g < n; g += (long long)gridDim.x * kBlock) {
    PokerTable<MAXP> t;
    uint32_t fb;
    packed_load(t, fb, S, g, n, pr.AP);
    uint32_t w[R];
    pkw_pack_row<MAXP, MT>(t, pr, 0.0f, w);
#pragma unroll
    for (int i = 0; i < R; ++i) wire[(size_t)g * R + i] = w[i];
  }
}

}  // namespace

struct PulsePokerHostEnv {
  PulsePokerConfig cfg;
  int device = 0, chunks = 1, R = 3, W = 16;
  uint64_t seed = 0, env_base = 0;
  uint32_t hand_counter = 0;
  bool first = true;
  long long N = 0, Nc = 0;                 // tables, tables per chunk (a multiple of 1024; the last chunk may be short)
  PulsePokerPacked st{};
  std::vector<void*> dev_allocs, pin_allocs;
  uint8_t* d_act = nullptr; uint32_t* d_off = nullptr; uint32_t* d_wire = nullptr;
  uint8_t* h_act = nullptr; uint32_t* h_off = nullptr; uint32_t* h_wire = nullptr;
  std::vector<uint32_t> live;              // bit b of live[w]: table 32 w + b was not done after the last call
  std::vector<long long> chunk_rows;       // live tables per chunk at the start of the next step
  std::vector<long long> last_rows;        // records per chunk of the last call (what sits in h_wire now)
  bool last_compacted = false;
  std::vector<cudaStream_t> streams;
  std::vector<cudaEvent_t> events;
  // "chunk c has landed": the kernel writes the call's epoch into d_flag[c], a 4-byte copy queued behind the chunk's
  // records brings it to h_flag[c] (pinned), the workers poll that word.  No driver thread is involved (stream host
  // callbacks cost 20-50 us each and are run one at a time by a thread the spinning workers can starve).
  uint32_t* d_flag = nullptr; volatile uint32_t* h_flag = nullptr;
  uint32_t epoch = 0;
  std::vector<uint32_t> prev_live;         // the live mask as it was when the last call began (= which tables have a record)
  std::atomic<int>* remaining = nullptr;   // per chunk: workers still expanding it (the last one rebuilds its offsets)
  pulse_host::Pool* pool = nullptr;
  long long last_h2d = 0, last_d2h = 0;
  // seconds, accumulated over calls: staging the actions, submitting GPU work, the expansion phase (wall), and summed
  // over workers: time spent waiting for records / time spent expanding
  double t_stage = 0, t_submit = 0, t_expand = 0, t_wait_sum = 0, t_work_sum = 0;
  std::vector<double> t_chunk_seen, t_chunk_done;   // diagnostic: seconds after the expansion phase began at which the LAST worker saw chunk c land / finished its slice of it, summed over calls
  long long calls = 0;
  ~PulsePokerHostEnv() {
    delete pool;
    delete[] remaining;
    cudaSetDevice(device);
    for (auto s : streams) cudaStreamDestroy(s);
    for (auto e : events) cudaEventDestroy(e);
    for (auto p : dev_allocs) cudaFree(p);
    for (auto p : pin_allocs) cudaFreeHost(p);
  }
};

namespace {

#define HE_CUDA(call)                                                       \
  do {                                                                      \
    cudaError_t e_ = (call);                                                \
    if (e_ != cudaSuccess) {                                                \
      pulse_set_error("%s: %s", #call, cudaGetErrorString(e_));             \
      return (int)e_;                                                       \
    }                                                                       \
  } while (0)

template <class T>
int dev_alloc(PulsePokerHostEnv* e, T** p, size_t count) {
  HE_CUDA(cudaMalloc(reinterpret_cast<void**>(p), count * sizeof(T) + 256));
  e->dev_allocs.push_back(*p);
  HE_CUDA(cudaMemset(*p, 0, count * sizeof(T)));
  return 0;
}
template <class T>
int pin_alloc(PulsePokerHostEnv* e, T** p, size_t count) {
  HE_CUDA(cudaMallocHost(reinterpret_cast<void**>(p), count * sizeof(T) + 256));
  e->pin_allocs.push_back(*p);
  return 0;
}

PokerParams params_of(const PulsePokerConfig& c) {
  PokerParams p; p.w1 = c.w1; p.w2 = c.w2; p.K = c.K; p.alpha = c.alpha;
  p.P = c.n_players; p.AP = c.active_players; p.M = c.max_players;
  return p;
}

#define HE_DISPATCH(P, M, CALL)                                                   \
  do {                                                                            \
    if ((P) <= 2 && (M) == 2) { constexpr int MP = 2, MT = 2; CALL; }             \
    else if ((P) <= 4 && (M) == 4) { constexpr int MP = 4, MT = 4; CALL; }        \
    else if ((P) <= 6 && (M) == 6) { constexpr int MP = 6, MT = 6; CALL; }        \
    else if ((P) <= 10 && (M) == 10) { constexpr int MP = 10, MT = 10; CALL; }    \
    else { pulse_set_error("host env: unsupported (n_players, max_players)"); return PULSE_EUNSUPPORTED; } \
  } while (0)

// per-warp offsets of one chunk for the NEXT step, and the chunk's record count
void rebuild_offsets(PulsePokerHostEnv* e, int c) {
  const long long first = c * e->Nc, last = (first + e->Nc < e->N) ? first + e->Nc : e->N;
  uint32_t* off = e->h_off + (first >> 5) + c;           // each chunk has (Nc / 32 + 1) entries
  uint32_t acc = 0;
  for (long long wg = first >> 5; wg < ((last + 31) >> 5); ++wg) {
    off[wg - (first >> 5)] = acc;
    acc += (uint32_t)__builtin_popcount(e->live[wg]);
  }
  off[((last + 31) >> 5) - (first >> 5)] = acc;
  e->chunk_rows[c] = acc;
}

double now_s() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

template <int MP, int MT>
void launch_step_t(PulsePokerHostEnv* e, const PokerParams& pr, long long first, long long cnt, const uint32_t* d_off,
                   uint32_t* d_rows, uint32_t* d_flag, cudaStream_t s) {
  constexpr int MINB = MP <= 2 ? 8 : MP <= 6 ? 6 : 4;   // resident CTAs per SM, as for the float32-row step
  poker_host_step_kernel<MP, MT, MINB><<<pulse_grid_for(cnt, kBlock), kBlock, 0, s>>>(
      e->st, pr, e->d_act, d_off, d_rows, e->cfg.starting_bbs, e->cfg.max_bbs, first, first + cnt, e->N, d_flag, e->epoch);
}

int launch_step(PulsePokerHostEnv* e, const PokerParams& pr, long long first, long long cnt, const uint32_t* d_off,
                uint32_t* d_rows, uint32_t* d_flag, cudaStream_t s) {
  HE_DISPATCH(pr.P, pr.M, (launch_step_t<MP, MT>(e, pr, first, cnt, d_off, d_rows, d_flag, s)));
  return pulse_check_launch("pulse_poker_host_step");
}

// Every worker walks the chunks in order: waits until the chunk's records have landed (its epoch word has arrived in
// pinned memory), expands its own slice, and the last worker out of a chunk rebuilds that chunk's offsets for the next
// step.  No barrier between chunks: chunk c+1's copy overlaps chunk c's expansion.  Worker 0 -- the calling thread --
// first runs `submit`, which queues the GPU work of every chunk: the others are already waiting for chunk 0 by then, so
// queueing chunks 1.. overlaps chunk 0's transfer instead of preceding the whole phase.
int expand_chunks(PulsePokerHostEnv* e, bool compacted, float* obs, float* reward, uint8_t* done, bool fresh,
                  const std::function<int()>& submit, bool records_only = false) {
  const int T = e->pool->size();
  for (int c = 0; c < e->chunks; ++c) {
    const long long first = c * e->Nc, last = (first + e->Nc < e->N) ? first + e->Nc : e->N;
    e->last_rows[c] = compacted ? e->chunk_rows[c] : last - first;
  }
  e->last_compacted = compacted;
  for (int c = 0; c < e->chunks; ++c) e->remaining[c].store(T, std::memory_order_relaxed);
  std::atomic<int> failed{0};
  std::vector<double> waits((size_t)T, 0.0), works((size_t)T, 0.0);
  const uint32_t epoch = e->epoch;
  const double t0 = now_s();
  e->pool->run([&, e, compacted, obs, reward, done, fresh, T, epoch](int wk) {
    if (wk == 0) {
      const double s0 = now_s();
      const int rc = submit();
      e->t_submit += now_s() - s0;
      if (rc) { failed.store(rc); return; }
    }
    for (int c = 0; c < e->chunks; ++c) {
      const double w0 = now_s();
      long long spins = 0;
      while (e->h_flag[c] != epoch) {
        if (failed.load(std::memory_order_relaxed)) return;
        if (++spins < 2048) _mm_pause();
        else std::this_thread::yield();
        if ((spins & 0xFFFF) == 0 && now_s() - w0 > 60.0) { failed.store(PULSE_ENOTREADY); return; }   // a dead GPU must not hang the caller
      }
      std::atomic_thread_fence(std::memory_order_acquire);
      const double w1 = now_s();
      if (wk == T - 1) e->t_chunk_seen[c] += w1 - t0;
      const long long first = c * e->Nc, last = (first + e->Nc < e->N) ? first + e->Nc : e->N;
      const long long per = (((last - first) + T - 1) / T + 31) & ~31LL;
      const long long g0 = first + wk * per, g1 = (g0 + per < last) ? g0 + per : last;
      if (g0 < g1) {
        PulseExpandArgs a;
        a.rows = e->h_wire + (size_t)first * e->R;
        a.off = compacted ? e->h_off + (first >> 5) + c : nullptr;
        a.live = e->live.data(); a.prev_live = e->prev_live.data(); a.chunk_first = first; a.g0 = g0; a.g1 = g1;
        a.row_words = e->R; a.obs_width = e->W; a.max_players = e->cfg.max_players;
        a.obs = obs; a.reward = reward; a.done = done; a.fresh = fresh; a.records_only = records_only;
        pulse_expand_range(a);
      }
      if (e->remaining[c].fetch_sub(1, std::memory_order_acq_rel) == 1) rebuild_offsets(e, c);
      const double w2 = now_s();
      if (wk == T - 1) e->t_chunk_done[c] += w2 - t0;
      waits[wk] += w1 - w0; works[wk] += w2 - w1;
    }
  });
  e->t_expand += now_s() - t0;
  for (int i = 0; i < T; ++i) { e->t_wait_sum += waits[i]; e->t_work_sum += works[i]; }
  const int f = failed.load();
  if (f == PULSE_ENOTREADY) pulse_set_error("host env: timed out waiting for the device");
  return f;
}

}  // namespace

extern "C" {

int pulse_poker_host_create(PulsePokerHostEnv** out, const PulsePokerConfig* cfg, const int32_t* host_hand_ranks,
                            size_t n_hand_ranks, uint64_t seed, uint64_t env_base, int device, int n_chunks,
                            int n_threads) {
  PULSE_REQUIRE(out && cfg, "null argument");
  *out = nullptr;
  PULSE_REQUIRE(cfg->n_players >= 2 && cfg->n_players <= 10, "n_players must be in 2..10");
  PULSE_REQUIRE(cfg->active_players >= 2 && cfg->active_players <= cfg->n_players, "active_players must be in 2..n_players");
  PULSE_REQUIRE(cfg->n_games > 0, "n_games must be positive");
  PULSE_REQUIRE((long long)cfg->n_players * cfg->max_bbs < 65536 && cfg->starting_bbs <= cfg->max_bbs,
                "n_players * max_bbs must be below 65536 for the packed records (use the device-buffer API otherwise)");
  PULSE_REQUIRE(host_hand_ranks && n_hand_ranks >= 36875, "the HandRanks table (at least its first 36,875 entries) is required");
  const int P = cfg->n_players, M = cfg->max_players;
  if (!((P <= 2 && M == 2) || (P > 2 && P <= 4 && M == 4) || (P > 4 && P <= 6 && M == 6) || (P > 6 && M == 10))) {
    pulse_set_error("host env supports max_players 2, 4, 6 or 10 with n_players in the matching bracket");
    return PULSE_EUNSUPPORTED;
  }
  PulsePokerHostEnv* e = new (std::nothrow) PulsePokerHostEnv();
  PULSE_REQUIRE(e != nullptr, "out of memory");
  e->cfg = *cfg; e->device = device; e->seed = seed; e->env_base = env_base;
  e->N = cfg->n_games; e->R = pkw_row_words(M); e->W = 13 + 3 * (M - 1);
  if (n_chunks < 1) n_chunks = 1;
  long long nc = ((e->N + n_chunks - 1) / n_chunks + 1023) & ~1023LL;
  e->Nc = nc; e->chunks = (int)((e->N + nc - 1) / nc);
  if (n_threads < 1) { n_threads = (int)std::thread::hardware_concurrency(); if (n_threads < 1) n_threads = 1; }
  // a worker hand-off costs a few microseconds: below ~16 K tables per worker the pool is slower than the caller's own
  // thread (256 tables, the reference's benchmark size: 87 us per step with 16 workers)
  if ((long long)n_threads > (e->N + 16383) / 16384) n_threads = (int)((e->N + 16383) / 16384);
  int rc = 0;
  auto fail = [&](int code) { delete e; return code; };
  if (cudaSetDevice(device) != cudaSuccess) { pulse_set_error("cudaSetDevice(%d) failed", device); return fail(PULSE_EINVAL); }
  const size_t N = (size_t)e->N, PN = N * P;
  int32_t* hr = nullptr;
  if ((rc = dev_alloc(e, &e->st.tbl, 4 * N)) || (rc = dev_alloc(e, &e->st.tbl2, 2 * N)) || (rc = dev_alloc(e, &e->st.stack, PN)) ||
      (rc = dev_alloc(e, &e->st.crb, PN)) || (rc = dev_alloc(e, &e->st.inv, PN)) || (rc = dev_alloc(e, &e->st.hole, PN)) ||
      (rc = dev_alloc(e, &e->st.ev, PN)) || (rc = dev_alloc(e, &e->st.hand_ctr, N)) || (rc = dev_alloc(e, &hr, 36875)) ||
      (rc = dev_alloc(e, &e->d_act, N)) || (rc = dev_alloc(e, &e->d_off, N / 32 + 2 * e->chunks + 2)) ||
      (rc = dev_alloc(e, &e->d_wire, N * e->R)) || (rc = pin_alloc(e, &e->h_act, N)) ||
      (rc = pin_alloc(e, &e->h_off, N / 32 + 2 * e->chunks + 2)) || (rc = pin_alloc(e, &e->h_wire, N * e->R)))
    return fail(rc);
  if (cudaMemcpy(hr, host_hand_ranks, 36875 * sizeof(int32_t), cudaMemcpyHostToDevice) != cudaSuccess) {
    pulse_set_error("uploading the HandRanks prefix failed"); return fail(PULSE_EINVAL);
  }
  e->st.hand_ranks = hr; e->st.n_hand_ranks = 36875; e->st.idx = nullptr; e->st.prev_invested = nullptr;
  e->live.assign(N / 32 + 2, 0u);
  e->chunk_rows.assign(e->chunks, 0);
  e->last_rows.assign(e->chunks, 0);
  for (int c = 0; c < e->chunks; ++c) {
    cudaStream_t s; cudaEvent_t ev;
    if (cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreateWithFlags(&ev, cudaEventDisableTiming) != cudaSuccess) {
      pulse_set_error("stream / event creation failed"); return fail(PULSE_EINVAL);
    }
    e->streams.push_back(s); e->events.push_back(ev);
  }
  e->pool = new pulse_host::Pool(n_threads);
  e->remaining = new std::atomic<int>[e->chunks];
  for (int c = 0; c < e->chunks; ++c) e->remaining[c].store(0);
  {
    uint32_t* hf = nullptr;
    if ((rc = dev_alloc(e, &e->d_flag, (size_t)e->chunks + 32)) || (rc = pin_alloc(e, &hf, (size_t)e->chunks + 32))) return fail(rc);
    for (int c = 0; c < e->chunks + 32; ++c) hf[c] = 0;
    e->h_flag = hf;
  }
  e->prev_live.assign(N / 32 + 2, 0u);
  e->t_chunk_seen.assign(e->chunks, 0.0); e->t_chunk_done.assign(e->chunks, 0.0);
  *out = e;
  return PULSE_OK;
}

void pulse_poker_host_destroy(PulsePokerHostEnv* env) { delete env; }

int pulse_poker_host_reset(PulsePokerHostEnv* e, int active_players, int rotation, float* obs_host, float* reward_host_or_null,
                           uint8_t* done_host_or_null) {
  PULSE_REQUIRE(e != nullptr, "null argument");   // obs_host == NULL: records only (see pulse_poker_host_step_records)
  PULSE_REQUIRE(active_players >= 2 && active_players <= e->cfg.n_players, "active_players must be in 2..n_players");
  HE_CUDA(cudaSetDevice(e->device));
  e->cfg.active_players = active_players;
  const PokerParams pr = params_of(e->cfg);
  e->epoch += 1;
  // reward / done scratch when the caller does not want them
  std::vector<float> rtmp; std::vector<uint8_t> dtmp;
  float* rw = reward_host_or_null; uint8_t* dn = done_host_or_null;
  if (!rw) { rtmp.resize((size_t)e->N); rw = rtmp.data(); }
  if (!dn) { dtmp.resize((size_t)e->N); dn = dtmp.data(); }
  // the deal (no float32 observation), then every table's record, copied chunk by chunk
  auto submit = [e, pr, rotation]() -> int {
    int rc = pulse_poker_packed_reset(&e->st, &e->cfg, nullptr, e->seed, e->env_base, e->hand_counter, e->first ? 0 : rotation,
                                      e->first ? 7 : 0, nullptr, e->streams[0]);
    if (rc) return rc;
    e->first = false;
    e->hand_counter += 1;
    const int grid = pulse_grid_for(e->N, kBlock);
    HE_DISPATCH(pr.P, pr.M, (poker_host_rows_kernel<MP, MT><<<grid, kBlock, 0, e->streams[0]>>>(e->st, pr, e->d_wire, e->N,
                                                                                              e->d_flag, e->chunks, e->epoch)));
    rc = pulse_check_launch("pulse_poker_host_reset");
    if (rc) return rc;
    HE_CUDA(cudaEventRecord(e->events[0], e->streams[0]));
    for (int c = 0; c < e->chunks; ++c) {
      const long long first = c * e->Nc, cnt = ((first + e->Nc < e->N) ? first + e->Nc : e->N) - first;
      if (c > 0) HE_CUDA(cudaStreamWaitEvent(e->streams[c], e->events[0], 0));
      HE_CUDA(cudaMemcpyAsync(e->h_wire + (size_t)first * e->R, e->d_wire + (size_t)first * e->R, (size_t)cnt * e->R * 4,
                              cudaMemcpyDeviceToHost, e->streams[c]));
      HE_CUDA(cudaMemcpyAsync(const_cast<uint32_t*>(e->h_flag) + c, e->d_flag + c, 4, cudaMemcpyDeviceToHost, e->streams[c]));
    }
    return 0;
  };
  e->last_h2d = 0; e->last_d2h = (long long)e->N * e->R * 4;
  return expand_chunks(e, false, obs_host, rw, dn, true, submit, obs_host == nullptr);
}

static int host_step_impl(PulsePokerHostEnv* e, const void* actions_host, int action_bytes, float* obs_host,
                          float* reward_host, uint8_t* done_host, bool records_only) {
  PULSE_REQUIRE(e && actions_host && done_host && (records_only || (obs_host && reward_host)), "null argument");
  PULSE_REQUIRE(action_bytes == 1 || action_bytes == 8, "action_bytes must be 1 (uint8) or 8 (int64)");
  PULSE_REQUIRE(!e->first, "reset first");
  HE_CUDA(cudaSetDevice(e->device));
  const PokerParams pr = params_of(e->cfg);
  // stage the actions in pinned memory as bytes -- unless the caller's byte array is page-locked already (the copy
  // engine can read it where it lies: 0.09 ms of a 2 ms step)
  const double t_stage0 = now_s();
  const uint8_t* act_src = e->h_act;
  if (action_bytes == 1) {
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, actions_host) == cudaSuccess && at.type == cudaMemoryTypeHost)
      act_src = static_cast<const uint8_t*>(actions_host);
    else
      (void)cudaGetLastError();   // an ordinary pageable pointer is not an error worth keeping
  }
  if (act_src == e->h_act) e->pool->run([=](int wk) {
    const int T = e->pool->size();
    const long long per = ((e->N + T - 1) / T + 63) & ~63LL;
    const long long a = wk * per, b = (a + per < e->N) ? a + per : e->N;
    if (action_bytes == 1) {
      if (a < b) memcpy(e->h_act + a, static_cast<const uint8_t*>(actions_host) + a, (size_t)(b - a));
    } else {
      const int64_t* src = static_cast<const int64_t*>(actions_host);
      // ids above 12 all behave like 13 (a raise of nothing, PokerGPU.py:255-273), negative ids match no branch (255)
      for (long long i = a; i < b; ++i) { const int64_t v = src[i]; e->h_act[i] = v < 0 ? (uint8_t)255 : v > 12 ? (uint8_t)13 : (uint8_t)v; }
    }
  });
  e->t_stage += now_s() - t_stage0;
  e->epoch += 1;
  long long h2d = 0, d2h = 0;
  for (int c = 0; c < e->chunks; ++c) {
    const long long first = c * e->Nc, cnt = ((first + e->Nc < e->N) ? first + e->Nc : e->N) - first;
    h2d += cnt + (((cnt + 31) >> 5) + 1) * 4;
    d2h += e->chunk_rows[c] * e->R * 4 + 4;
  }
  e->last_h2d = h2d; e->last_d2h = d2h; e->calls += 1;
  auto submit = [e, pr, act_src]() -> int {
    for (int c = 0; c < e->chunks; ++c) {
      const long long first = c * e->Nc, cnt = ((first + e->Nc < e->N) ? first + e->Nc : e->N) - first;
      const long long n_off = ((cnt + 31) >> 5) + 1;
      cudaStream_t s = e->streams[c];
      uint32_t* d_off = e->d_off + (first >> 5) + c;
      HE_CUDA(cudaMemcpyAsync(e->d_act + first, act_src + first, (size_t)cnt, cudaMemcpyHostToDevice, s));
      HE_CUDA(cudaMemcpyAsync(d_off, e->h_off + (first >> 5) + c, (size_t)n_off * 4, cudaMemcpyHostToDevice, s));
      uint32_t* d_rows = e->d_wire + (size_t)first * e->R;
      int rc = launch_step(e, pr, first, cnt, d_off, d_rows, e->d_flag + c, s);
      if (rc) return rc;
      const size_t bytes = (size_t)e->chunk_rows[c] * e->R * 4;
      if (bytes) HE_CUDA(cudaMemcpyAsync(e->h_wire + (size_t)first * e->R, d_rows, bytes, cudaMemcpyDeviceToHost, s));
      HE_CUDA(cudaMemcpyAsync(const_cast<uint32_t*>(e->h_flag) + c, e->d_flag + c, 4, cudaMemcpyDeviceToHost, s));
    }
    return 0;
  };
  return expand_chunks(e, true, obs_host, reward_host, done_host, false, submit, records_only);
}

int pulse_poker_host_step(PulsePokerHostEnv* e, const void* actions_host, int action_bytes, float* obs_host,
                          float* reward_host, uint8_t* done_host) {
  return host_step_impl(e, actions_host, action_bytes, obs_host, reward_host, done_host, false);
}

// The step without the float32 expansion: the packed records stay in the env's pinned buffer (pulse_poker_host_records),
// only the done flags are written out.  For consumers that read the records directly (or expand rows on demand with
// pulse_poker_wire_expand_host); the host then touches ~4x less memory per step.
int pulse_poker_host_step_records(PulsePokerHostEnv* e, const void* actions_host, int action_bytes, uint8_t* done_host) {
  return host_step_impl(e, actions_host, action_bytes, nullptr, nullptr, done_host, true);
}

// chunk c of the last call: its records (in table order), how many, and the table range the chunk covers; the tables
// that have a record are the set bits of pulse_poker_host_live_mask() within that range (all of them after a reset)
int pulse_poker_host_records(const PulsePokerHostEnv* e, int chunk, const uint32_t** rows, long long* n_rows,
                             long long* first_table, long long* n_tables) {
  PULSE_REQUIRE(e && chunk >= 0 && chunk < e->chunks, "bad chunk");
  const long long first = (long long)chunk * e->Nc, last = (first + e->Nc < e->N) ? first + e->Nc : e->N;
  if (rows) *rows = e->h_wire + (size_t)first * e->R;
  if (n_rows) *n_rows = e->last_rows[chunk];
  if (first_table) *first_table = first;
  if (n_tables) *n_tables = last - first;
  return PULSE_OK;
}

// bit b of word w: table 32 w + b was live when the last call began, i.e. it has a record in that call's output
const uint32_t* pulse_poker_host_live_mask(const PulsePokerHostEnv* e) { return e ? e->prev_live.data() : nullptr; }

int pulse_poker_host_info(const PulsePokerHostEnv* e, int* row_bytes, int* obs_width, int* chunks, int* threads,
                          long long* last_h2d_bytes, long long* last_d2h_bytes) {
  PULSE_REQUIRE(e != nullptr, "null env");
  if (row_bytes) *row_bytes = e->R * 4;
  if (obs_width) *obs_width = e->W;
  if (chunks) *chunks = e->chunks;
  if (threads) *threads = e->pool->size();
  if (last_h2d_bytes) *last_h2d_bytes = e->last_h2d;
  if (last_d2h_bytes) *last_d2h_bytes = e->last_d2h;
  return PULSE_OK;
}

// seconds accumulated since creation (or the last call with reset != 0): out6 = {calls, staging the actions, submitting
// GPU work, the expansion phase (wall clock), worker time waiting for records (summed over workers), worker time
// expanding (summed over workers)}
int pulse_poker_host_timing(PulsePokerHostEnv* e, double* out6, int reset) {
  PULSE_REQUIRE(e && out6, "null argument");
  out6[0] = (double)e->calls; out6[1] = e->t_stage; out6[2] = e->t_submit; out6[3] = e->t_expand;
  out6[4] = e->t_wait_sum; out6[5] = e->t_work_sum;
  if (reset) {
    e->calls = 0; e->t_stage = e->t_submit = e->t_expand = e->t_wait_sum = e->t_work_sum = 0;
    for (auto& v : e->t_chunk_seen) v = 0;
    for (auto& v : e->t_chunk_done) v = 0;
  }
  return PULSE_OK;
}

// diagnostic: per chunk, seconds (summed over the calls since the last timing reset, resets included) after the
// expansion phase began at which the last worker saw the chunk's records land (out[2 c]) and finished its slice of it
// (out[2 c + 1]); n_chunks entries are written, at most max_chunks
int pulse_poker_host_chunk_timing(const PulsePokerHostEnv* e, double* out, int max_chunks) {
  PULSE_REQUIRE(e && out, "null argument");
  for (int c = 0; c < e->chunks && c < max_chunks; ++c) { out[2 * c] = e->t_chunk_seen[c]; out[2 * c + 1] = e->t_chunk_done[c]; }
  return e->chunks;
}

// pure host function (no GPU): one record -> the reference's float32 row; exported for tests and for callers that
// keep the compact records and expand rows on demand
int pulse_poker_wire_expand_host(const uint32_t* rows, long long n_rows, int max_players, int vectorised,
                                 float* obs_out, float* reward_out, uint8_t* done_out) {
  PULSE_REQUIRE(rows && obs_out && reward_out && done_out, "null argument");
  PULSE_REQUIRE(max_players >= 2 && max_players <= 16, "bad table size");
  const int R = pkw_row_words(max_players), W = 13 + 3 * (max_players - 1);
  if (!vectorised) {   // the portable definition (poker_wire.cuh)
    for (long long i = 0; i < n_rows; ++i)
      pkw_expand_row(rows + (size_t)i * R, max_players, obs_out + (size_t)i * W, reward_out + i, done_out + i);
    return PULSE_OK;
  }
  // the host env's own expander (AVX2 when the CPU has it), driven as one chunk in which every table has a record
  std::vector<uint32_t> live((size_t)(n_rows + 31) / 32 + 1, 0u);
  PulseExpandArgs a;
  std::vector<uint32_t> prev(live.size(), 0u);
  a.rows = rows; a.off = nullptr; a.live = live.data(); a.prev_live = prev.data(); a.chunk_first = 0; a.g0 = 0; a.g1 = n_rows;
  a.row_words = R; a.obs_width = W; a.max_players = max_players;
  a.obs = obs_out; a.reward = reward_out; a.done = done_out; a.fresh = true; a.records_only = false;
  pulse_expand_range(a);
  return PULSE_OK;
}

// pure host function: one step of the host env's bookkeeping without a GPU -- `rows` holds the records of the tables
// whose bit is set in live_inout (compacted in table order; every table when fresh != 0); obs / reward / done are the
// caller's persistent arrays of n_tables entries; live_inout / prev_live_inout (ceil(n / 32) words) are updated exactly as
// the env updates them.  Exists so the CPU suite can drive the expander through whole hands.
int pulse_poker_wire_expand_step_host(const uint32_t* rows, uint32_t* live_inout, uint32_t* prev_live_inout,
                                      long long n_tables, int max_players, int fresh, float* obs, float* reward,
                                      uint8_t* done) {
  PULSE_REQUIRE(rows && live_inout && prev_live_inout && obs && reward && done, "null argument");
  PULSE_REQUIRE(max_players >= 2 && max_players <= 16 && n_tables >= 0, "bad arguments");
  const long long groups = (n_tables + 31) >> 5;
  std::vector<uint32_t> off((size_t)groups + 1, 0u);
  uint32_t acc = 0;
  for (long long w = 0; w < groups; ++w) { off[w] = acc; acc += (uint32_t)__builtin_popcount(live_inout[w]); }
  PulseExpandArgs a;
  a.rows = rows; a.off = fresh ? nullptr : off.data(); a.live = live_inout; a.prev_live = prev_live_inout;
  a.chunk_first = 0; a.g0 = 0; a.g1 = n_tables;
  a.row_words = pkw_row_words(max_players); a.obs_width = 13 + 3 * (max_players - 1); a.max_players = max_players;
  a.obs = obs; a.reward = reward; a.done = done; a.fresh = fresh != 0; a.records_only = false;
  pulse_expand_range(a);
  return PULSE_OK;
}

const char* pulse_poker_host_expand_isa(void) { return pulse_expand_isa(); }
}
