// Measures what bounds the end-to-end (host-buffer) path of the env engine on a GPU box: pinned H2D / D2H copy
// bandwidth per GPU and across all visible GPUs at once, and the host's streaming-store bandwidth per thread count
// (the observation expander writes fp32 rows with non-temporal stores).  Evidence for DESIGN.md section 5; not product.
//   nvcc -O3 -std=c++17 -o tools/_build/hostlink_probe tools/hostlink_probe.cu && tools/_build/hostlink_probe
#include <cuda_runtime.h>
#include <emmintrin.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <chrono>
#include <thread>
#include <vector>

static double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

struct Dev { int id; char* d; char* h_in; char* h_out; cudaStream_t s_in, s_out; };

static void copy_test(std::vector<Dev>& devs, size_t bytes, int reps, bool h2d, bool d2h, const char* label) {
  for (auto& v : devs) { CK(cudaSetDevice(v.id)); CK(cudaDeviceSynchronize()); }
  double t0 = now();
  for (int r = 0; r < reps; ++r)
    for (auto& v : devs) {
      CK(cudaSetDevice(v.id));
      if (h2d) CK(cudaMemcpyAsync(v.d, v.h_in, bytes, cudaMemcpyHostToDevice, v.s_in));
      if (d2h) CK(cudaMemcpyAsync(v.h_out, v.d + bytes, bytes, cudaMemcpyDeviceToHost, v.s_out));
    }
  for (auto& v : devs) { CK(cudaSetDevice(v.id)); CK(cudaDeviceSynchronize()); }
  double dt = now() - t0;
  double gb = (double)bytes * reps * devs.size() * ((h2d ? 1 : 0) + (d2h ? 1 : 0)) / 1e9;
  printf("{\"test\": \"%s\", \"gpus\": %zu, \"block_mb\": %.0f, \"aggregate_gbs\": %.1f, \"per_gpu_gbs\": %.1f}\n", label, devs.size(),
         bytes / 1e6, gb / dt, gb / dt / devs.size());
}

// streaming fill: each thread writes its slice with 16 B non-temporal stores
static void fill_nt(char* p, size_t bytes, int reps) {
  __m128i v = _mm_set1_epi32(0x3f800000);
  for (int r = 0; r < reps; ++r) {
    for (size_t i = 0; i < bytes; i += 64) {
      _mm_stream_si128((__m128i*)(p + i), v); _mm_stream_si128((__m128i*)(p + i + 16), v);
      _mm_stream_si128((__m128i*)(p + i + 32), v); _mm_stream_si128((__m128i*)(p + i + 48), v);
    }
    _mm_sfence();
  }
}
static void fill_plain(char* p, size_t bytes, int reps) {
  for (int r = 0; r < reps; ++r) { memset(p, r & 255, bytes); }
}
// read 24 B records, write 64 B rows (the expander's shape)
static void expand_like(const char* src, char* dst, size_t rows, int reps) {
  for (int r = 0; r < reps; ++r) {
    for (size_t i = 0; i < rows; ++i) {
      const uint64_t* q = (const uint64_t*)(src + i * 24);
      __m128i a = _mm_set_epi64x((long long)q[1], (long long)q[0]), b = _mm_set1_epi64x((long long)q[2]);
      char* o = dst + i * 64;
      _mm_stream_si128((__m128i*)o, a); _mm_stream_si128((__m128i*)(o + 16), b);
      _mm_stream_si128((__m128i*)(o + 32), a); _mm_stream_si128((__m128i*)(o + 48), b);
    }
    _mm_sfence();
  }
}

int main(int argc, char** argv) {
  int n = 0;
  CK(cudaGetDeviceCount(&n));
  int hw = (int)std::thread::hardware_concurrency();
  printf("{\"test\": \"info\", \"gpus\": %d, \"host_threads\": %d}\n", n, hw);
  const size_t B = 256u << 20;
  std::vector<Dev> devs;
  for (int i = 0; i < n; ++i) {
    Dev v; v.id = i;
    CK(cudaSetDevice(i));
    CK(cudaMalloc(&v.d, 2 * B)); CK(cudaMallocHost(&v.h_in, B)); CK(cudaMallocHost(&v.h_out, B));
    memset(v.h_in, 1, B); memset(v.h_out, 2, B);
    CK(cudaStreamCreate(&v.s_in)); CK(cudaStreamCreate(&v.s_out));
    devs.push_back(v);
  }
  for (int k = 1; k <= n; k *= 2) {
    std::vector<Dev> sub(devs.begin(), devs.begin() + k);
    for (size_t bytes : {(size_t)(4u << 20), (size_t)(64u << 20), B}) {
      copy_test(sub, bytes, 2, false, true, "warm");
      copy_test(sub, bytes, (int)((size_t)(4096u << 20) / bytes / 4 + 4), false, true, "d2h");
      copy_test(sub, bytes, (int)((size_t)(4096u << 20) / bytes / 4 + 4), true, false, "h2d");
      copy_test(sub, bytes, (int)((size_t)(4096u << 20) / bytes / 4 + 4), true, true, "h2d+d2h");
    }
  }
  // host streaming-store bandwidth (pageable memory, first touched by the writing thread)
  const size_t H = 1024u << 20;
  char* buf = (char*)aligned_alloc(4096, H);
  char* src = (char*)aligned_alloc(4096, H / 2);
  memset(buf, 0, H); memset(src, 3, H / 2);
  for (int mode = 0; mode < 3; ++mode)
    for (int t = 1; t <= hw; t *= 2) {
      std::vector<std::thread> th;
      const int reps = 4;
      double t0 = now();
      for (int k = 0; k < t; ++k) {
        size_t sl = H / t / 4096 * 4096;
        if (mode == 0) th.emplace_back(fill_nt, buf + k * sl, sl, reps);
        else if (mode == 1) th.emplace_back(fill_plain, buf + k * sl, sl, reps);
        else { size_t rows = sl / 64; th.emplace_back(expand_like, src + (size_t)k * (rows * 24), buf + k * sl, rows, reps); }
      }
      for (auto& x : th) x.join();
      double dt = now() - t0;
      printf("{\"test\": \"%s\", \"threads\": %d, \"write_gbs\": %.1f}\n", mode == 0 ? "host_nt_store" : mode == 1 ? "host_memset" : "host_expand_24B_to_64B",
             t, (double)(H / t / 4096 * 4096) * t * reps / dt / 1e9);
    }
  // the same expander writing into PINNED memory while all GPUs stream D2H into other pinned buffers
  return 0;
}
