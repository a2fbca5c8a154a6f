// 2048 with HOST buffers: the end-to-end path of the batched 2048 env (include/pulse_envs.h, "host-buffer 2048").
//
// The reference's step hands back an int32 board of tile VALUES per env plus reward and done
// (/root/reference/environments/2048/TFE.py:152-189, board :121): 69 B per board-step, which is what bounded the end
// to end rate of the device-buffer class with host arrays (7.6e8 board-steps/s = the PCIe link,
// profiles/r2j_bench_lines.jsonl).  The engine's own board is 8 B (sixteen 4-bit exponents), so that is what crosses
// PCIe -- 1 B of action in, 8 B board + 4 B reward + 1 B done out -- and a pool of host threads rebuilds the int32 tile
// values in the caller's array with streaming stores (csrc/host_expand.cpp).  The batch is cut into chunks, each on its
// own stream: chunk c + 1's step and copies run while the threads expand chunk c.  Chance is keyed by global board id,
// so the chunked launches draw exactly what one whole-batch launch of pulse_tfe_step draws.
//
// The handle owns its device state, pinned staging buffers, streams and threads; the caller's arrays may be pageable.
// One handle = one env on one GPU; calls on one handle are serial.
#include "common.cuh"
#include "host_expand.h"
#include "host_pool.h"

#include <emmintrin.h>
#include <string.h>

#include <atomic>
#include <chrono>
#include <new>

struct PulseTfeHostEnv {
  int device = 0, side = 4, chunks = 1, auto_reset = 1;
  long long N = 0, Nc = 0;
  uint64_t seed = 0, env_base = 0;
  uint32_t t = 0;                 // tape counter: one per reset / step
  bool started = false;
  uint64_t* d_boards = nullptr; int64_t* d_score = nullptr; uint8_t* d_done = nullptr; uint8_t* d_act = nullptr;
  int32_t* d_reward = nullptr;
  uint64_t* h_boards = nullptr; uint8_t* h_act = nullptr; int32_t* h_reward = nullptr; uint8_t* h_done = nullptr;
  // "chunk c has landed": the call's epoch goes to a device word at the head of the chunk's stream work and comes back
  // with a 4-byte copy queued behind the chunk's device-to-host copies; the workers poll it in pinned memory.  (Both
  // hops are real DMA copies: a host-to-host cudaMemcpyAsync is carried out by the calling thread once the stream has
  // drained, which would turn the submission loop into a wait -- profiles/r2z_zerocopy_probe.jsonl, "dma_chain".)
  volatile uint32_t* h_flag = nullptr; uint32_t* h_epoch = nullptr; uint32_t* d_flag = nullptr;
  uint32_t epoch = 0;
  std::vector<cudaStream_t> streams;
  std::vector<void*> dev_allocs, pin_allocs;
  pulse_host::Pool* pool = nullptr;
  long long last_h2d = 0, last_d2h = 0;
  ~PulseTfeHostEnv() {
    delete pool;
    cudaSetDevice(device);
    for (auto s : streams) cudaStreamDestroy(s);
    for (auto p : dev_allocs) cudaFree(p);
    for (auto p : pin_allocs) cudaFreeHost(p);
  }
};

namespace {

#define TH_CUDA(call)                                                       \
  do {                                                                      \
    cudaError_t e_ = (call);                                                \
    if (e_ != cudaSuccess) {                                                \
      pulse_set_error("%s: %s", #call, cudaGetErrorString(e_));             \
      return (int)e_;                                                       \
    }                                                                       \
  } while (0)

template <class T>
int dev_alloc(PulseTfeHostEnv* e, T** p, size_t count) {
  TH_CUDA(cudaMalloc(reinterpret_cast<void**>(p), count * sizeof(T) + 256));
  e->dev_allocs.push_back(*p);
  TH_CUDA(cudaMemset(*p, 0, count * sizeof(T)));
  return 0;
}
template <class T>
int pin_alloc(PulseTfeHostEnv* e, T** p, size_t count) {
  TH_CUDA(cudaMallocHost(reinterpret_cast<void**>(p), count * sizeof(T) + 256));
  e->pin_allocs.push_back(*p);
  memset(*p, 0, count * sizeof(T));
  return 0;
}

double now_s() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

// Every worker walks the chunks in order: waits for the chunk's flag, rebuilds its slice of the observations and copies
// its slice of rewards / done flags.  Worker 0 -- the calling thread -- first queues the GPU work of every chunk.
int run_chunks(PulseTfeHostEnv* e, int32_t* obs, int32_t* reward, uint8_t* done, const std::function<int()>& submit) {
  const int T = e->pool->size();
  std::atomic<int> failed{0};
  const uint32_t epoch = e->epoch;
  e->pool->run([&, e, obs, reward, done, T, epoch](int wk) {
    if (wk == 0) {
      const int rc = submit();
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
      const long long first = c * e->Nc, last = (first + e->Nc < e->N) ? first + e->Nc : e->N;
      const long long per = (((last - first) + T - 1) / T + 31) & ~31LL;
      const long long g0 = first + wk * per, g1 = (g0 + per < last) ? g0 + per : last;
      if (g0 < g1) {
        if (obs) pulse_tfe_expand_range(e->h_boards, obs, g0, g1, e->side);
        if (reward) memcpy(reward + g0, e->h_reward + g0, (size_t)(g1 - g0) * sizeof(int32_t));
        if (done) memcpy(done + g0, e->h_done + g0, (size_t)(g1 - g0));
      }
    }
  });
  const int f = failed.load();
  if (f == PULSE_ENOTREADY) pulse_set_error("2048 host env: timed out waiting for the device");
  return f;
}

}  // namespace

extern "C" {

int pulse_tfe_host_create(PulseTfeHostEnv** out, long long n_envs, int side, int auto_reset, uint64_t seed,
                          uint64_t env_base, int device, int n_chunks, int n_threads) {
  PULSE_REQUIRE(out != nullptr, "null argument");
  *out = nullptr;
  PULSE_REQUIRE(n_envs > 0 && n_envs <= 0x7FFFFFFFLL, "n_envs must be in 1..2^31-1");
  PULSE_REQUIRE(side >= 2 && side <= 4, "side must be 2, 3 or 4");
  PulseTfeHostEnv* e = new (std::nothrow) PulseTfeHostEnv();
  PULSE_REQUIRE(e != nullptr, "out of memory");
  e->device = device; e->side = side; e->auto_reset = auto_reset ? 1 : 0; e->seed = seed; e->env_base = env_base; e->N = n_envs;
  if (n_chunks < 1) n_chunks = 1;
  const long long nc = ((e->N + n_chunks - 1) / n_chunks + 1023) & ~1023LL;
  e->Nc = nc; e->chunks = (int)((e->N + nc - 1) / nc);
  if (n_threads < 1) { n_threads = (int)std::thread::hardware_concurrency(); if (n_threads < 1) n_threads = 1; }
  if ((long long)n_threads > (e->N + 16383) / 16384) n_threads = (int)((e->N + 16383) / 16384);   // see host_env.cu
  auto fail = [&](int code) { delete e; return code; };
  if (cudaSetDevice(device) != cudaSuccess) { pulse_set_error("cudaSetDevice(%d) failed", device); return fail(PULSE_EINVAL); }
  const size_t N = (size_t)e->N;
  int rc = 0;
  uint32_t* hf = nullptr;
  if ((rc = dev_alloc(e, &e->d_boards, N)) || (rc = dev_alloc(e, &e->d_score, N)) || (rc = dev_alloc(e, &e->d_done, N)) ||
      (rc = dev_alloc(e, &e->d_act, N)) || (rc = dev_alloc(e, &e->d_reward, N)) || (rc = pin_alloc(e, &e->h_boards, N)) ||
      (rc = pin_alloc(e, &e->h_act, N)) || (rc = pin_alloc(e, &e->h_reward, N)) || (rc = pin_alloc(e, &e->h_done, N)) ||
      (rc = pin_alloc(e, &hf, (size_t)e->chunks + 32)) || (rc = pin_alloc(e, &e->h_epoch, 16)) ||
      (rc = dev_alloc(e, &e->d_flag, (size_t)e->chunks + 32)))
    return fail(rc);
  e->h_flag = hf;
  for (int c = 0; c < e->chunks; ++c) {
    cudaStream_t s;
    if (cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking) != cudaSuccess) { pulse_set_error("stream creation failed"); return fail(PULSE_EINVAL); }
    e->streams.push_back(s);
  }
  if ((rc = pulse_tfe_init(e->streams[0])) != 0) return fail(rc);    // this device's row-move table
  if (cudaStreamSynchronize(e->streams[0]) != cudaSuccess) { pulse_set_error("building the row table failed"); return fail(PULSE_EINVAL); }
  e->pool = new pulse_host::Pool(n_threads);
  *out = e;
  return PULSE_OK;
}

void pulse_tfe_host_destroy(PulseTfeHostEnv* env) { delete env; }

/* TFE.reset for every board (TFE.py:143-149): obs_host int32 [N, side, side] (or NULL), done_host_or_null cleared. */
int pulse_tfe_host_reset(PulseTfeHostEnv* e, int32_t* obs_host, uint8_t* done_host_or_null) {
  PULSE_REQUIRE(e != nullptr, "null argument");
  TH_CUDA(cudaSetDevice(e->device));
  e->epoch += 1;
  *e->h_epoch = e->epoch;
  auto submit = [e]() -> int {
    int rc = pulse_tfe_reset(e->d_boards, e->d_score, e->d_done, nullptr, e->seed, e->env_base, e->t, nullptr, (int)e->N,
                             e->side, e->streams[0]);
    if (rc) return rc;
    cudaStream_t s = e->streams[0];
    TH_CUDA(cudaMemcpyAsync(e->d_flag, e->h_epoch, 4, cudaMemcpyHostToDevice, s));
    TH_CUDA(cudaMemcpyAsync(e->h_boards, e->d_boards, (size_t)e->N * 8, cudaMemcpyDeviceToHost, s));
    TH_CUDA(cudaMemsetAsync(e->d_reward, 0, (size_t)e->N * 4, s));
    for (int c = 0; c < e->chunks; ++c)
      TH_CUDA(cudaMemcpyAsync(const_cast<uint32_t*>(e->h_flag) + c, e->d_flag, 4, cudaMemcpyDeviceToHost, s));
    return 0;
  };
  memset(e->h_reward, 0, (size_t)e->N * 4);
  memset(e->h_done, 0, (size_t)e->N);
  e->last_h2d = 0; e->last_d2h = e->N * 8;
  const int rc = run_chunks(e, obs_host, nullptr, done_host_or_null, submit);
  if (rc == 0) { e->t += 1; e->started = true; }
  return rc;
}

/* TFE.step for every board (TFE.py:152-189; auto_reset as pulse_tfe_step).  actions_host uint8 or int64 [N];
 * obs_host int32 [N, side, side] (or NULL: boards only), reward_host int32 [N], done_host uint8 [N]. */
int pulse_tfe_host_step(PulseTfeHostEnv* e, const void* actions_host, int action_bytes, int32_t* obs_host,
                        int32_t* reward_host, uint8_t* done_host) {
  PULSE_REQUIRE(e && actions_host && reward_host && done_host, "null argument");
  PULSE_REQUIRE(action_bytes == 1 || action_bytes == 8, "action_bytes must be 1 (uint8) or 8 (int64)");
  PULSE_REQUIRE(e->started, "reset first");
  TH_CUDA(cudaSetDevice(e->device));
  e->pool->run([=](int wk) {           // the actions as bytes in pinned memory
    const int T = e->pool->size();
    const long long per = ((e->N + T - 1) / T + 63) & ~63LL;
    const long long a = wk * per, b = (a + per < e->N) ? a + per : e->N;
    if (a >= b) return;
    if (action_bytes == 1) memcpy(e->h_act + a, static_cast<const uint8_t*>(actions_host) + a, (size_t)(b - a));
    else {
      const int64_t* src = static_cast<const int64_t*>(actions_host);
      for (long long i = a; i < b; ++i) e->h_act[i] = (uint8_t)(src[i] & 3);   // as the kernel reads an int64 action
    }
  });
  e->epoch += 1;
  *e->h_epoch = e->epoch;
  auto submit = [e]() -> int {
    for (int c = 0; c < e->chunks; ++c) {
      const long long first = c * e->Nc, cnt = ((first + e->Nc < e->N) ? first + e->Nc : e->N) - first;
      cudaStream_t s = e->streams[c];
      TH_CUDA(cudaMemcpyAsync(e->d_flag + c, e->h_epoch, 4, cudaMemcpyHostToDevice, s));
      TH_CUDA(cudaMemcpyAsync(e->d_act + first, e->h_act + first, (size_t)cnt, cudaMemcpyHostToDevice, s));
      const int rc = pulse_tfe_step(e->d_boards + first, e->d_act + first, 1, nullptr, nullptr, e->seed, e->env_base + (uint64_t)first,
                                    e->t, nullptr, e->auto_reset, nullptr, e->d_reward + first, e->d_done + first,
                                    e->d_score + first, (int)cnt, e->side, s);
      if (rc) return rc;
      TH_CUDA(cudaMemcpyAsync(e->h_boards + first, e->d_boards + first, (size_t)cnt * 8, cudaMemcpyDeviceToHost, s));
      TH_CUDA(cudaMemcpyAsync(e->h_reward + first, e->d_reward + first, (size_t)cnt * 4, cudaMemcpyDeviceToHost, s));
      TH_CUDA(cudaMemcpyAsync(e->h_done + first, e->d_done + first, (size_t)cnt, cudaMemcpyDeviceToHost, s));
      TH_CUDA(cudaMemcpyAsync(const_cast<uint32_t*>(e->h_flag) + c, e->d_flag + c, 4, cudaMemcpyDeviceToHost, s));
    }
    return 0;
  };
  e->last_h2d = e->N; e->last_d2h = e->N * 13;
  const int rc = run_chunks(e, obs_host, reward_host, done_host, submit);
  if (rc == 0) e->t += 1;
  return rc;
}

/* the packed boards of the last call, in pinned host memory (for callers that keep the 8-byte form) */
const uint64_t* pulse_tfe_host_boards(const PulseTfeHostEnv* e) { return e ? e->h_boards : nullptr; }

/* total_score of every board (TFE.total_score, :183): a device-to-host copy, outside the step path */
int pulse_tfe_host_scores(PulseTfeHostEnv* e, int64_t* score_host) {
  PULSE_REQUIRE(e && score_host, "null argument");
  TH_CUDA(cudaSetDevice(e->device));
  TH_CUDA(cudaMemcpy(score_host, e->d_score, (size_t)e->N * 8, cudaMemcpyDeviceToHost));
  return PULSE_OK;
}

int pulse_tfe_host_info(const PulseTfeHostEnv* e, int* chunks, int* threads, long long* last_h2d_bytes,
                        long long* last_d2h_bytes) {
  PULSE_REQUIRE(e != nullptr, "null env");
  if (chunks) *chunks = e->chunks;
  if (threads) *threads = e->pool->size();
  if (last_h2d_bytes) *last_h2d_bytes = e->last_h2d;
  if (last_d2h_bytes) *last_d2h_bytes = e->last_d2h;
  return PULSE_OK;
}

/* pure host function: packed boards -> int32 tile values (the expander the env uses; exported for tests) */
int pulse_tfe_expand_host(const uint64_t* boards, long long n, int side, int32_t* obs_out) {
  PULSE_REQUIRE(boards && obs_out && n >= 0 && side >= 2 && side <= 4, "bad arguments");
  pulse_tfe_expand_range(boards, obs_out, 0, n, side);
  return PULSE_OK;
}
}
