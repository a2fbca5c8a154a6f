// Tetris kernels and their C ABI (include/pulse_envs.h).  Rules: tetris_body.cuh / DESIGN.md section 6.
// State is rows uint16 [N][20] row-major (40 B per env; this tensor IS the observation, as the
// reference's TFE returns its board array itself), piece uint8 [N], lines int32 [N], done uint8 [N].
// One thread per env; each block brings its contiguous [kBlock][20] slab of rows into shared memory
// with 16 B coalesced loads (prefetched one slab ahead), threads drop / lock / clear on their own 20 rows there (dynamic row
// indexing stays out of local memory), and the slab goes back with 16 B coalesced stores.  The 28-entry
// shape table is staged in shared memory too.  HBM-bound: DESIGN.md 4.4.
#include "tetris_body.cuh"

namespace {
constexpr int kBlock = 128;
constexpr int kSlabWords = kBlock * TET_H / 2;  // uint32 words in a full slab

__constant__ uint32_t c_shapes[TET_SHAPES] = TET_SHAPE_TABLE;

// A slab is the kBlock envs' rows: kBlock * 40 B = 320 uint4.  Thread t moves uint4 t, t + 128 and (t < 64)
// t + 256; a partial last slab falls back to 4-byte words.
struct SlabRegs { uint4 v[3]; };

__device__ __forceinline__ void slab_load(SlabRegs& r, const uint32_t* __restrict__ g, int words) {
  if (words == kSlabWords) {
    const uint4* g4 = reinterpret_cast<const uint4*>(g);
    r.v[0] = g4[threadIdx.x]; r.v[1] = g4[threadIdx.x + kBlock];
    if (threadIdx.x < kSlabWords / 4 - 2 * kBlock) r.v[2] = g4[threadIdx.x + 2 * kBlock];
  }
}
__device__ __forceinline__ void slab_to_smem(uint32_t* s, const SlabRegs& r, const uint32_t* __restrict__ g, int words) {
  if (words == kSlabWords) {
    uint4* s4 = reinterpret_cast<uint4*>(s);
    s4[threadIdx.x] = r.v[0]; s4[threadIdx.x + kBlock] = r.v[1];
    if (threadIdx.x < kSlabWords / 4 - 2 * kBlock) s4[threadIdx.x + 2 * kBlock] = r.v[2];
  } else {
    for (int i = threadIdx.x; i < words; i += kBlock) s[i] = g[i];
  }
}
__device__ __forceinline__ void slab_store(uint32_t* __restrict__ g, const uint32_t* s, int words) {
  if (words == kSlabWords) {
    const uint4* s4 = reinterpret_cast<const uint4*>(s);
    uint4* g4 = reinterpret_cast<uint4*>(g);
    g4[threadIdx.x] = s4[threadIdx.x]; g4[threadIdx.x + kBlock] = s4[threadIdx.x + kBlock];
    if (threadIdx.x < kSlabWords / 4 - 2 * kBlock) g4[threadIdx.x + 2 * kBlock] = s4[threadIdx.x + 2 * kBlock];
  } else {
    for (int i = threadIdx.x; i < words; i += kBlock) g[i] = s[i];
  }
}

// Per-env scalars of a slab, as loaded: byte fields stay raw 32-bit registers (ld.global.u8 zero-extends, nothing is
// masked, tested or widened here) so that no instruction consumes a prefetched value before the iteration that
// uses it -- a `done[g] != 0` or `& 0xff` right behind the load parks the warp on the DRAM round trip the
// prefetch exists to hide.
struct TetScalars { uint32_t pc, dn, a8; int nx, rp; int32_t ln; long long a64; };

__device__ __forceinline__ uint32_t tet_ld_u8(const uint8_t* p) {
  uint32_t v;
  asm volatile("ld.global.u8 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
}

template <typename ActT>
__device__ __forceinline__ TetScalars tet_load_scalars(const uint8_t* __restrict__ piece, const int32_t* __restrict__ lines,
                                                       const uint8_t* __restrict__ done, const ActT* __restrict__ actions,
                                                       const uint8_t* __restrict__ next_piece,
                                                       const uint8_t* __restrict__ reset_piece, long long g, bool valid) {
  TetScalars s;
  s.pc = 0; s.nx = -1; s.rp = -1; s.ln = 0; s.dn = 0; s.a8 = 0; s.a64 = 0;
  if (valid) {
    s.pc = tet_ld_u8(piece + g); s.ln = lines[g]; s.dn = tet_ld_u8(done + g);
    if (sizeof(ActT) == 1) s.a8 = tet_ld_u8(reinterpret_cast<const uint8_t*>(actions) + g);
    else s.a64 = (long long)actions[g];
    if (next_piece) s.nx = next_piece[g];
    if (reset_piece) s.rp = reset_piece[g];
  }
  return s;
}

// Persistent CTAs (at most kCtasPerSm per SM) walk the slabs with a grid stride.  The next slab and its
// per-env scalars are requested into registers before the current slab is processed, so the HBM round
// trip of slab i+1 hides behind the work on slab i; two shared-memory slabs alternate so one barrier per
// slab is enough between "threads finished reading their rows" and "next slab lands".
constexpr int kCtasPerSm = 12;

template <typename ActT>
__global__ void __launch_bounds__(kBlock)
tetris_step_kernel(uint16_t* __restrict__ rows, uint8_t* __restrict__ piece, int32_t* __restrict__ lines,
                   uint8_t* __restrict__ done, const ActT* __restrict__ actions, const uint8_t* __restrict__ next_piece,
                   const uint8_t* __restrict__ reset_piece, uint64_t seed, uint64_t env_base, uint32_t counter,
                   uint32_t* counter_dev, int auto_reset, int32_t* __restrict__ reward_out, long long n) {
  __shared__ __align__(16) uint32_t slab[2][kSlabWords];
  __shared__ uint32_t shapes[TET_SHAPES];
  const long long stride = (long long)gridDim.x * kBlock;
  long long base = (long long)blockIdx.x * kBlock;
  // first slab and scalars go out before anything waits on the device counter
  SlabRegs regs;
  TetScalars sc;
  int cnt = 0;
  if (base < n) {
    cnt = (int)((n - base) < kBlock ? (n - base) : kBlock);
    slab_load(regs, reinterpret_cast<const uint32_t*>(rows + base * TET_H), cnt * TET_H / 2);
    sc = tet_load_scalars(piece, lines, done, actions, next_piece, reset_piece, base + threadIdx.x, threadIdx.x < cnt);
  }
  if (threadIdx.x < TET_SHAPES) shapes[threadIdx.x] = c_shapes[threadIdx.x];
  const uint32_t dev_ctr = pulse_counter_read(counter_dev);
  counter += dev_ctr;
  int buf = 0;
  for (; base < n; base += stride, buf ^= 1) {
    uint32_t* gslab = reinterpret_cast<uint32_t*>(rows + base * TET_H);
    const int words = cnt * TET_H / 2;
    slab_to_smem(slab[buf], regs, gslab, words);
    const TetScalars cur = sc;
    const bool valid = threadIdx.x < cnt;
    // request the next slab
    const long long nbase = base + stride;
    int ncnt = 0;
    if (nbase < n) {
      ncnt = (int)((n - nbase) < kBlock ? (n - nbase) : kBlock);
      slab_load(regs, reinterpret_cast<const uint32_t*>(rows + nbase * TET_H), ncnt * TET_H / 2);
      sc = tet_load_scalars(piece, lines, done, actions, next_piece, reset_piece, nbase + threadIdx.x, threadIdx.x < ncnt);
    }
    const long long g = base + threadIdx.x;
    int nx = cur.nx, rp = cur.rp;
    if (valid && (nx < 0 || rp < 0)) {
      const PulseU4 blk = pulse_tape(seed, env_base + (uint64_t)g, PULSE_STREAM_TETRIS_PIECE, counter);
      if (nx < 0) nx = (int)pulse_mulhi32(blk.x, 7u);
      if (rp < 0) rp = (int)pulse_mulhi32(blk.y, 7u);
    }
    __syncthreads();                                   // slab[buf] complete (and shapes, first time round)
    if (valid) {
      const int pc0 = (int)(cur.pc & 0xFFu);
      const bool dn0 = (cur.dn & 0xFFu) != 0;
      const long long act = sizeof(ActT) == 1 ? (long long)(cur.a8 & 0xFFu) : cur.a64;
      int pc = pc0; int32_t ln = cur.ln; bool dn = dn0;
      int32_t reward;
      tet_step_one(reinterpret_cast<uint16_t*>(slab[buf]) + threadIdx.x * TET_H, shapes, pc, ln, dn, act, nx, rp,
                   auto_reset, reward);
      if (pc != pc0) piece[g] = (uint8_t)pc;
      if (ln != cur.ln) lines[g] = ln;
      if (dn != dn0) done[g] = dn;
      reward_out[g] = reward;
    }
    __syncthreads();                                   // every thread's rows are final
    slab_store(gslab, slab[buf], words);
    cnt = ncnt;
  }
  pulse_counter_advance(counter_dev, dev_ctr);
}

__global__ void __launch_bounds__(256)
tetris_reset_kernel(uint16_t* __restrict__ rows, uint8_t* __restrict__ piece, int32_t* __restrict__ lines,
                    uint8_t* __restrict__ done, const uint8_t* __restrict__ first_piece, uint64_t seed,
                    uint64_t env_base, uint32_t counter, long long n) {
  uint32_t* w = reinterpret_cast<uint32_t*>(rows);
  const long long words = n * (TET_H / 2), stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < words; i += stride) w[i] = 0;
  for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < n; g += stride) {
    piece[g] = first_piece ? first_piece[g]
                           : (uint8_t)pulse_mulhi32(pulse_tape(seed, env_base + (uint64_t)g, PULSE_STREAM_TETRIS_PIECE, counter).y, 7u);
    lines[g] = 0; done[g] = 0;
  }
}
}  // namespace

extern "C" {

int pulse_tetris_reset(uint16_t* rows, uint8_t* piece, int32_t* lines, uint8_t* done, const uint8_t* first_piece_or_null,
                       uint64_t seed, uint64_t env_base, uint32_t counter, int n_envs, cudaStream_t stream) {
  PULSE_REQUIRE(n_envs >= 0, "n_envs < 0");
  if (n_envs == 0) return PULSE_OK;
  PULSE_REQUIRE(rows && piece && lines && done, "null buffer");
  tetris_reset_kernel<<<pulse_grid_for((long long)n_envs * 10, 256), 256, 0, stream>>>(
      rows, piece, lines, done, first_piece_or_null, seed, env_base, counter, (long long)n_envs);
  return pulse_check_launch("pulse_tetris_reset");
}

int pulse_tetris_step(uint16_t* rows, uint8_t* piece, int32_t* lines, uint8_t* done, const void* actions,
                      int action_bytes, const uint8_t* next_piece_or_null, const uint8_t* reset_piece_or_null,
                      uint64_t seed, uint64_t env_base, uint32_t counter, uint32_t* counter_dev_or_null,
                      int auto_reset, int32_t* reward_out, int n_envs, cudaStream_t stream) {
  PULSE_REQUIRE(n_envs >= 0, "n_envs < 0");
  PULSE_REQUIRE(action_bytes == 1 || action_bytes == 8, "actions must be uint8 or int64");
  if (n_envs == 0) return PULSE_OK;
  PULSE_REQUIRE(rows && piece && lines && done && actions && reward_out, "null buffer");
  PULSE_REQUIRE((reinterpret_cast<uintptr_t>(rows) & 15u) == 0, "rows must be 16-byte aligned");
  const long long need = ((long long)n_envs + kBlock - 1) / kBlock;
  const long long cap = (long long)pulse_sm_count() * kCtasPerSm;
  const int grid = (int)(need < cap ? need : cap);  // persistent CTAs
  if (action_bytes == 1)
    tetris_step_kernel<uint8_t><<<grid, kBlock, 0, stream>>>(rows, piece, lines, done, (const uint8_t*)actions,
                                                             next_piece_or_null, reset_piece_or_null, seed, env_base,
                                                             counter, counter_dev_or_null, auto_reset, reward_out,
                                                             (long long)n_envs);
  else
    tetris_step_kernel<int64_t><<<grid, kBlock, 0, stream>>>(rows, piece, lines, done, (const int64_t*)actions,
                                                             next_piece_or_null, reset_piece_or_null, seed, env_base,
                                                             counter, counter_dev_or_null, auto_reset, reward_out,
                                                             (long long)n_envs);
  return pulse_check_launch("pulse_tetris_step");
}
}
