// Shared device/host helpers for libpulse_envs: error plumbing, launch sizing, and the
// counter-based Philox-4x32-10 chance tape.
//
// Chance tape: every random draw an env consumes is
//   philox4x32_10(counter = {env_id_lo, env_id_hi, draw_counter, stream_tag},
//                 key     = {seed_lo, seed_hi})
// so results depend only on (seed, global env id, counter) and never on grid shape,
// GPU count or launch order.  oracle/oracle_core.c holds an independent restatement.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "pulse_envs.h"

// stream tags of the chance tape
#define PULSE_STREAM_POKER_DECK 1u
#define PULSE_STREAM_ACTION 2u
#define PULSE_STREAM_TFE_SPAWN 3u
#define PULSE_STREAM_TFE_RESET 4u
#define PULSE_STREAM_BJ_DECK 5u
#define PULSE_STREAM_P2D_INIT 6u
#define PULSE_STREAM_TETRIS_PIECE 7u
#define PULSE_STREAM_EQUITY 8u   // | sample index << 4 (poker_packed_body.cuh, showdown equity)

void pulse_set_error(const char* fmt, ...);
int pulse_check_launch(const char* what);
int pulse_sm_count(void);   // multiprocessors of the current device (cudaDevAttrMultiProcessorCount, cached); 148 on B200

#define PULSE_REQUIRE(cond, msg)                      \
  do {                                                \
    if (!(cond)) {                                    \
      pulse_set_error("%s: %s", __func__, msg);       \
      return PULSE_EINVAL;                            \
    }                                                 \
  } while (0)

struct PulseU4 { uint32_t x, y, z, w; };

struct PulseI2 { int32_t x, y; };
__host__ __device__ __forceinline__ PulseI2 pulse_ld_int2(const int32_t* p) {  // 8-byte aligned pair
#ifdef __CUDA_ARCH__
  int2 v = *reinterpret_cast<const int2*>(p);
  PulseI2 r; r.x = v.x; r.y = v.y; return r;
#else
  PulseI2 r; r.x = p[0]; r.y = p[1]; return r;
#endif
}

// Per-env logic is written as __host__ __device__ functions so that tests/hostemu can compile the
// very same code for the CPU and check it against the oracle without a GPU.  The product library
// only ever launches the __global__ wrappers.
#define PULSE_HD __host__ __device__ __forceinline__
#ifdef __CUDA_ARCH__
#define PULSE_POPC(x) __popc(x)
#define PULSE_FFS(x) __ffs(x)
#define PULSE_FMUL(a, b) __fmul_rn(a, b)
#define PULSE_FADD(a, b) __fadd_rn(a, b)
#define PULSE_FSUB(a, b) __fsub_rn(a, b)
#define PULSE_FDIV(a, b) __fdiv_rn(a, b)
#define PULSE_FSQRT(a) __fsqrt_rn(a)
#define PULSE_F2I_RZ(a) __float2int_rz(a)
#define PULSE_LDG(p) __ldg(p)
#else
#include <math.h>
#define PULSE_POPC(x) __builtin_popcount(x)
#define PULSE_FFS(x) __builtin_ffs(x)
#define PULSE_FMUL(a, b) ((a) * (b))
#define PULSE_FADD(a, b) ((a) + (b))
#define PULSE_FSUB(a, b) ((a) - (b))
#define PULSE_FDIV(a, b) ((a) / (b))
#define PULSE_FSQRT(a) sqrtf(a)
#define PULSE_F2I_RZ(a) ((int)(a))
#define PULSE_LDG(p) (*(p))
#endif

__host__ __device__ __forceinline__ uint32_t pulse_mulhi32(uint32_t a, uint32_t b) {
#ifdef __CUDA_ARCH__
  return __umulhi(a, b);
#else
  return (uint32_t)(((uint64_t)a * (uint64_t)b) >> 32);
#endif
}

__host__ __device__ __forceinline__ PulseU4 pulse_philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2,
                                                                 uint32_t c3, uint32_t k0, uint32_t k1) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    uint32_t hi0 = pulse_mulhi32(M0, c0), lo0 = M0 * c0;
    uint32_t hi1 = pulse_mulhi32(M1, c2), lo1 = M1 * c2;
    uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
    c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  PulseU4 o; o.x = c0; o.y = c1; o.z = c2; o.w = c3;
  return o;
}

// One 4-draw block of the tape.
__host__ __device__ __forceinline__ PulseU4 pulse_tape(uint64_t seed, uint64_t env_id, uint32_t stream,
                                                        uint32_t counter) {
  return pulse_philox4x32_10((uint32_t)env_id, (uint32_t)(env_id >> 32), counter, stream,
                             (uint32_t)seed, (uint32_t)(seed >> 32));
}

// Device-resident step counter (so a step launch can be captured once in a CUDA graph and replayed with
// an advancing chance-tape counter): ctr[0] = steps taken, ctr[1] = ticket of finished blocks.  Every
// block reads ctr[0] when it starts; the last block to finish bumps it for the next launch.  A block
// can only be last after every other block has finished, hence after every block has read.
#ifdef __CUDACC__
__device__ __forceinline__ uint32_t pulse_counter_read(uint32_t* ctr) {
  return ctr ? *reinterpret_cast<volatile uint32_t*>(ctr) : 0u;
}
__device__ __forceinline__ void pulse_counter_advance(uint32_t* ctr, uint32_t seen) {
  if (!ctr) return;
  __syncthreads();
  if (threadIdx.x == 0) {
    // no fence needed: `seen` was consumed (a true dependency) before this point, and the ticket is an atomic
    if (atomicAdd(ctr + 1, 1u) == gridDim.x - 1) {
      ctr[1] = 0;
      *reinterpret_cast<volatile uint32_t*>(ctr) = seen + 1u;
    }
  }
}
#endif

// TMA bulk copy global -> shared, completion on an mbarrier (sm_90+; SASS: UBLKCP).  Used to stage
// lookup tables: one elected thread issues the copies, nobody spends registers or issue slots on them.
#ifdef __CUDACC__
__device__ __forceinline__ uint32_t pulse_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void pulse_mbar_init(uint64_t* bar, uint32_t arrivals) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(pulse_smem_u32(bar)), "r"(arrivals) : "memory");
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void pulse_mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(pulse_smem_u32(bar)), "r"(bytes) : "memory");
}
// bytes: multiple of 16; dst and src 16-byte aligned
__device__ __forceinline__ void pulse_bulk_g2s(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(pulse_smem_u32(smem_dst)), "l"(gmem_src), "r"(bytes), "r"(pulse_smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void pulse_mbar_wait(uint64_t* bar, uint32_t phase) {
  asm volatile(
      "{\n"
      ".reg .pred P1;\n"
      "PULSE_MBAR_WAIT:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
      "@P1 bra PULSE_MBAR_DONE;\n"
      "bra PULSE_MBAR_WAIT;\n"
      "PULSE_MBAR_DONE:\n"
      "}" ::"r"(pulse_smem_u32(bar)), "r"(phase) : "memory");
}
#endif

// Grid sizing: grids are a whole number of waves (SM count x resident CTAs; 148 SMs on B200) where the problem is
// big enough, otherwise just enough CTAs to cover it.
static inline int pulse_grid_for(long long n, int block, int ctas_per_sm = 8) {
  long long need = (n + block - 1) / block;
  long long wave = (long long)pulse_sm_count() * ctas_per_sm;
  if (need <= wave) return (int)(need > 0 ? need : 1);
  long long waves = (need + wave - 1) / wave;
  return (int)(waves * wave);  // grid-stride loops absorb the tail
}
