// Probe for the host-buffer path (csrc/host_env.cu): how do a DMA copy chain and a kernel writing straight into mapped
// pinned host memory compare -- bandwidth, first-data latency, and both again while host threads stream stores into
// memory (what the expander does).  One JSON line per test.
//   nvcc -O2 -gencode arch=compute_100a,code=sm_100a -o tools/_build/zerocopy_probe tools/zerocopy_probe.cu -lpthread
#include <cuda_runtime.h>
#include <immintrin.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <atomic>
#include <chrono>
#include <thread>
#include <vector>

static double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("{\"error\": \"%s: %s\"}\n", #x, cudaGetErrorString(e_)); return 1; } } while (0)

// every thread writes `words_per_thread` consecutive 16-byte words of a contiguous region: a warp covers 512 B per store
__global__ void write_host16(uint4* dst, size_t n16, uint32_t tag) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x)
    dst[i] = make_uint4(tag, (uint32_t)i, tag, (uint32_t)i);
}
// 4-byte stores, a warp covers 128 B per store
__global__ void write_host4(uint32_t* dst, size_t n4, uint32_t tag) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) dst[i] = tag + (uint32_t)i;
}
// chunk kernel as the env would run it: write the chunk, fence, count finished blocks, the last one raises the flag
__global__ void write_chunk_flag(uint32_t* dst, size_t n4, uint32_t tag, unsigned* counter, volatile uint32_t* flag, uint32_t epoch) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) dst[i] = tag + (uint32_t)i;
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    if (atomicAdd(counter, 1u) == gridDim.x - 1) { *counter = 0; __threadfence_system(); *flag = epoch; }
  }
}
__global__ void fill_dev(uint32_t* p, size_t n) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) p[i] = (uint32_t)i;
}

static std::atomic<bool> g_stop{false};
static void nt_load(char* p, size_t bytes) {   // streaming stores until told to stop
  while (!g_stop.load(std::memory_order_relaxed)) {
    for (size_t i = 0; i + 32 <= bytes; i += 32) _mm256_stream_si256((__m256i*)(p + i), _mm256_set1_epi32(7));
    _mm_sfence();
  }
}

int main(int argc, char** argv) {
  setvbuf(stdout, nullptr, _IOLBF, 0);
  const size_t total = 50u << 20, chunks = 8, chunk = total / chunks;   // what one 4M-table heads-up step ships
  uint32_t *h = nullptr, *d = nullptr, *hflag = nullptr, *hsrc = nullptr;
  unsigned* dcount = nullptr;
  CK(cudaSetDevice(0));
  CK(cudaMallocHost((void**)&h, total + 4096));
  CK(cudaMallocHost((void**)&hflag, 4096));
  CK(cudaMallocHost((void**)&hsrc, 4096));
  CK(cudaMalloc((void**)&d, total));
  CK(cudaMalloc((void**)&dcount, 64 * sizeof(unsigned)));
  CK(cudaMemset(dcount, 0, 64 * sizeof(unsigned)));
  fill_dev<<<1184, 256>>>(d, total / 4);
  CK(cudaDeviceSynchronize());
  std::vector<cudaStream_t> st(chunks);
  for (auto& s : st) CK(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
  const int hw = (int)std::thread::hardware_concurrency();
  const size_t load_bytes = (size_t)256 << 20;
  char* loadbuf = nullptr;
  if (posix_memalign((void**)&loadbuf, 4096, load_bytes) != 0) { printf("{\"error\": \"posix_memalign\"}\n"); return 1; }
  memset(loadbuf, 0, load_bytes);
  for (int loaded = 0; loaded < 2; ++loaded) {
    std::vector<std::thread> th;
    g_stop = false;
    const int nload = loaded ? (hw > 4 ? hw - 2 : 2) : 0;
    const size_t slice = nload ? (load_bytes / nload) & ~(size_t)4095 : 0;
    for (int i = 0; i < nload; ++i) th.emplace_back(nt_load, loadbuf + (size_t)i * slice, slice);
    if (loaded) std::this_thread::sleep_for(std::chrono::milliseconds(50));
    // A: DMA chain per chunk (records copy + 4-byte flag copy), chunks on their own streams, as host_env.cu does
    for (int rep = 0; rep < 5; ++rep) {
      const uint32_t epoch = 100 + rep + 10 * loaded;
      *hsrc = epoch;
      volatile uint32_t* f = hflag;
      const double t0 = now();
      for (size_t c = 0; c < chunks; ++c) {
        CK(cudaMemcpyAsync((char*)h + c * chunk, (char*)d + c * chunk, chunk, cudaMemcpyDeviceToHost, st[c]));
        CK(cudaMemcpyAsync(hflag + c, hsrc, 4, cudaMemcpyHostToHost, st[c]));
      }
      const double t1 = now();
      double first = 0, last = 0;
      for (size_t c = 0; c < chunks; ++c) { while (f[c] != epoch) _mm_pause(); if (c == 0) first = now(); }
      last = now();
      if (rep == 4)
        printf("{\"test\": \"dma_chain\", \"host_store_threads\": %d, \"submit_ms\": %.3f, \"first_chunk_ms\": %.3f, \"all_chunks_ms\": %.3f, \"gbs\": %.1f}\n",
               nload, (t1 - t0) * 1e3, (first - t0) * 1e3, (last - t0) * 1e3, total / (last - t0) / 1e9);
    }
    // B: one kernel per chunk writing straight into pinned host memory, last block raises the chunk's flag
    for (int w = 0; w < 2; ++w) {
      for (int rep = 0; rep < 5; ++rep) {
        const uint32_t epoch = 200 + rep + 10 * loaded + 100 * w;
        volatile uint32_t* f = hflag;
        const double t0 = now();
        for (size_t c = 0; c < chunks; ++c)
          write_chunk_flag<<<w == 0 ? 148 : 592, 256, 0, st[c]>>>((uint32_t*)((char*)h + c * chunk), chunk / 4, epoch, dcount + c, hflag + c, epoch);
        const double t1 = now();
        double first = 0, last = 0;
        for (size_t c = 0; c < chunks; ++c) { while (f[c] != epoch) _mm_pause(); if (c == 0) first = now(); }
        last = now();
        if (rep == 4)
          printf("{\"test\": \"zero_copy_chunks\", \"blocks_per_chunk\": %d, \"host_store_threads\": %d, \"submit_ms\": %.3f, \"first_chunk_ms\": %.3f, \"all_chunks_ms\": %.3f, \"gbs\": %.1f}\n",
                 w == 0 ? 148 : 592, nload, (t1 - t0) * 1e3, (first - t0) * 1e3, (last - t0) * 1e3, total / (last - t0) / 1e9);
      }
    }
    // C: plain bandwidth of kernel stores into host memory, 16-byte and 4-byte per thread
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    for (int v = 0; v < 2; ++v) {
      float best = 1e9f;
      for (int rep = 0; rep < 4; ++rep) {
        CK(cudaEventRecord(e0));
        if (v == 0) write_host16<<<1184, 256>>>((uint4*)h, total / 16, rep);
        else write_host4<<<1184, 256>>>(h, total / 4, rep);
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        if (ms < best) best = ms;
      }
      printf("{\"test\": \"zero_copy_bandwidth\", \"bytes_per_thread_store\": %d, \"host_store_threads\": %d, \"ms\": %.3f, \"gbs\": %.1f}\n",
             v == 0 ? 16 : 4, nload, best, total / (best * 1e-3) / 1e9);
    }
    // D: plain D2H copy of the whole 50 MB
    {
      float best = 1e9f;
      for (int rep = 0; rep < 4; ++rep) {
        CK(cudaEventRecord(e0));
        CK(cudaMemcpyAsync(h, d, total, cudaMemcpyDeviceToHost, 0));
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        if (ms < best) best = ms;
      }
      printf("{\"test\": \"dma_d2h\", \"host_store_threads\": %d, \"ms\": %.3f, \"gbs\": %.1f}\n", nload, best, total / (best * 1e-3) / 1e9);
    }
    g_stop = true;
    for (auto& t : th) t.join();
  }
  return 0;
}
