// Host expander of the PokerGPU host-buffer path (csrc/host_env.cu): packed step records -> the reference's float32
// observation rows, rewards and done flags in the caller's arrays.  The record (poker_wire.cuh) is laid out so that
// this is "widen and convert": eight int8 columns, then uint16 columns in observation order.  Heads-up rows are 64 B =
// one cache line and go out as two 32-byte streaming stores (no read-for-ownership of a line that is fully rewritten).
// Plain host C++ (no CUDA): compiled for the baseline x86-64 ISA, with an AVX2 body selected at run time.
#include "host_expand.h"

#include <immintrin.h>
#include <string.h>

namespace {

inline void expand_row_portable(const uint32_t* w, int W, int R, float* o, float* reward, uint8_t* done) {
  const int8_t* b = reinterpret_cast<const int8_t*>(w);
  for (int i = 0; i < 8; ++i) o[i] = (float)b[i];
  const uint16_t* u = reinterpret_cast<const uint16_t*>(w + 2);
  for (int p = 0; p < W - 8; ++p) o[8 + p] = (float)u[p];
  o[12] = (float)(u[4] & 0x7FFFu);
  *done = (uint8_t)(u[4] >> 15);
  memcpy(reward, w + R - 1, 4);
}

__attribute__((target("avx2"))) inline void expand_row_avx2(const uint32_t* w, int W, int R, float* o, float* reward,
                                                            uint8_t* done, bool line_rows, bool stream16) {
  const __m256i keep = _mm256_setr_epi32(0xFFFF, 0xFFFF, 0xFFFF, 0xFFFF, 0x7FFF, 0xFFFF, 0xFFFF, 0xFFFF);  // column 12: status without the done bit
  const __m256 f0 = _mm256_cvtepi32_ps(_mm256_cvtepi8_epi32(_mm_loadl_epi64(reinterpret_cast<const __m128i*>(w))));
  const __m256 f1 = _mm256_cvtepi32_ps(_mm256_and_si256(
      _mm256_cvtepu16_epi32(_mm_loadu_si128(reinterpret_cast<const __m128i*>(w + 2))), keep));
  if (line_rows) {           // heads-up, 32-byte aligned rows: the whole cache line, streamed
    _mm256_stream_ps(o, f0);
    _mm256_stream_ps(o + 8, f1);
  } else if (stream16) {     // wider rows (W a multiple of 4, 16-byte aligned): streamed in 16-byte pieces
    _mm_stream_ps(o, _mm256_castps256_ps128(f0)); _mm_stream_ps(o + 4, _mm256_extractf128_ps(f0, 1));
    _mm_stream_ps(o + 8, _mm256_castps256_ps128(f1)); _mm_stream_ps(o + 12, _mm256_extractf128_ps(f1, 1));
    const uint16_t* u = reinterpret_cast<const uint16_t*>(w + 2);
    for (int p = 8; p < W - 8; p += 4)
      _mm_stream_ps(o + 8 + p, _mm_cvtepi32_ps(_mm_cvtepu16_epi32(_mm_loadl_epi64(reinterpret_cast<const __m128i*>(u + p)))));
  } else {
    _mm256_storeu_ps(o, f0);
    _mm256_storeu_ps(o + 8, f1);
    const uint16_t* u = reinterpret_cast<const uint16_t*>(w + 2);
    int p = 8;
    for (; p + 8 <= W - 8; p += 8)
      _mm256_storeu_ps(o + 8 + p, _mm256_cvtepi32_ps(_mm256_cvtepu16_epi32(_mm_loadu_si128(reinterpret_cast<const __m128i*>(u + p)))));
    for (; p < W - 8; ++p) o[8 + p] = (float)u[p];
  }
  *done = (uint8_t)((w[4] >> 15) & 1u);
  memcpy(reward, w + R - 1, 4);
}

// One 32-table group at a time.  Rewards and done flags of the WHOLE group are rebuilt in two small local arrays --
// a table that had already finished: reward 0 (PokerGPU.py:632), done 1, its observation row stays as it is -- and
// written out as whole lines (STORE_GROUP), so the reward / done arrays are never read: storing only the live tables'
// entries made every line of both arrays a read-for-ownership plus a write-back (40 MB of the ~190 MB a 4M-table
// step moves through host memory).
#define EXPAND_BODY(ROW_CALL, STORE_GROUP)                                                                      \
  const int R = a.row_words, W = a.obs_width;                                                                   \
  for (long long wg = a.g0 >> 5; wg < ((a.g1 + 31) >> 5); ++wg) {                                               \
    const long long gb = wg << 5;                                                                               \
    const int cnt = (int)((a.g1 - gb) < 32 ? (a.g1 - gb) : 32);                                                 \
    const uint32_t mask = a.fresh ? 0xFFFFFFFFu : a.live[wg];                                                   \
    uint32_t next = 0;                                                                                          \
    size_t slot = a.off ? a.off[wg - (a.chunk_first >> 5)] : (size_t)(gb - a.chunk_first);                      \
    const uint32_t in_range = cnt == 32 ? 0xFFFFFFFFu : ((1u << cnt) - 1u);   /* the batch's last group may be short */ \
    a.prev_live[wg] = mask;                                                                                      \
    alignas(32) float rw[32];                                                                                    \
    alignas(32) uint8_t dn[32];                                                                                  \
    for (int i = 0; i < 32; ++i) { rw[i] = 0.0f; dn[i] = 1; }                                                    \
    uint32_t m = mask & in_range;                                                                               \
    while (m) {                                                                                                 \
      const int b = __builtin_ctz(m);                                                                           \
      m &= m - 1;                                                                                               \
      const long long g = gb + b;                                                                               \
      const uint32_t* w = a.rows + slot * R;                                                                    \
      ++slot;                                                                                                   \
      uint8_t d;                                                                                                \
      ROW_CALL;                                                                                                 \
      dn[b] = d;                                                                                                \
      next |= (uint32_t)(d ? 0u : 1u) << b;                                                                     \
    }                                                                                                           \
    STORE_GROUP;                                                                                                \
    a.live[wg] = next;                                                                                          \
  }

#define STORE_GROUP_PLAIN                                                            \
  do {                                                                               \
    memcpy(a.reward + gb, rw, (size_t)cnt * sizeof(float));                          \
    memcpy(a.done + gb, dn, (size_t)cnt);                                            \
  } while (0)

void expand_range_portable(const PulseExpandArgs& a) {
  EXPAND_BODY(expand_row_portable(w, W, R, a.obs + g * W, rw + b, &d), STORE_GROUP_PLAIN)
}

__attribute__((target("avx2"))) void expand_range_avx2(const PulseExpandArgs& a) {
  const bool line_rows = a.obs_width == 16 && (reinterpret_cast<uintptr_t>(a.obs) & 31u) == 0;
  const bool stream16 = !line_rows && (a.obs_width & 3) == 0 && (reinterpret_cast<uintptr_t>(a.obs) & 15u) == 0;
  // whole groups on 32-byte aligned arrays: four streamed 32-byte pieces of rewards, one of done flags
  const bool stream_rd = (reinterpret_cast<uintptr_t>(a.reward) & 31u) == 0 && (reinterpret_cast<uintptr_t>(a.done) & 31u) == 0;
  EXPAND_BODY(expand_row_avx2(w, W, R, a.obs + g * W, rw + b, &d, line_rows, stream16),
              if (stream_rd && cnt == 32) {
                for (int i = 0; i < 4; ++i) _mm256_stream_ps(a.reward + gb + 8 * i, _mm256_load_ps(rw + 8 * i));
                _mm256_stream_si256(reinterpret_cast<__m256i*>(a.done + gb), _mm256_load_si256(reinterpret_cast<const __m256i*>(dn)));
              } else STORE_GROUP_PLAIN)
  _mm_sfence();
}

bool have_avx2() {
  static const bool v = __builtin_cpu_supports("avx2");
  return v;
}

}  // namespace

// records-only mode: the done flag of every record (bit 15 of the status column) into the caller's done array and
// the live mask; nothing is expanded
void scan_range_records(const PulseExpandArgs& a) {
  const int R = a.row_words;
  for (long long wg = a.g0 >> 5; wg < ((a.g1 + 31) >> 5); ++wg) {
    const long long gb = wg << 5;
    const int cnt = (int)((a.g1 - gb) < 32 ? (a.g1 - gb) : 32);
    const uint32_t in_range = cnt == 32 ? 0xFFFFFFFFu : ((1u << cnt) - 1u);
    const uint32_t mask = a.fresh ? 0xFFFFFFFFu : a.live[wg];
    size_t slot = a.off ? a.off[wg - (a.chunk_first >> 5)] : (size_t)(gb - a.chunk_first);
    uint32_t next = 0, m = mask & in_range;
    a.prev_live[wg] = mask;
    while (m) {
      const int b = __builtin_ctz(m);
      m &= m - 1;
      const uint8_t d = (uint8_t)((a.rows[slot * R + 4] >> 15) & 1u);
      ++slot;
      a.done[gb + b] = d;
      next |= (uint32_t)(d ? 0u : 1u) << b;
    }
    a.live[wg] = next;
  }
}

void pulse_expand_range(const PulseExpandArgs& a) {
  if (a.records_only) { scan_range_records(a); return; }
  if (have_avx2()) expand_range_avx2(a);
  else expand_range_portable(a);
}

const char* pulse_expand_isa(void) { return have_avx2() ? "avx2" : "portable"; }

// ---- 2048: packed board -> int32 tile values ---------------------------------------------------------------------
namespace {

inline void tfe_expand_portable(const uint64_t* boards, int32_t* obs, long long g0, long long g1, int side) {
  for (long long g = g0; g < g1; ++g) {
    const uint64_t b = boards[g];
    int32_t* o = obs + g * side * side;
    for (int r = 0; r < side; ++r)
      for (int c = 0; c < side; ++c) {
        const unsigned e = (unsigned)((b >> (16 * r + 4 * c)) & 15u);
        o[r * side + c] = e ? (int32_t)(1u << e) : 0;
      }
  }
}

// 4 x 4 boards, 32-byte aligned rows: nibbles -> bytes, two byte shuffles look up the low and high byte of 1 << e
// (0 for e = 0), widened to int32 and streamed as one 64-byte line
__attribute__((target("avx2"))) void tfe_expand_avx2_side4(const uint64_t* boards, int32_t* obs, long long g0, long long g1) {
  const __m128i lut_lo = _mm_setr_epi8(0, 2, 4, 8, 16, 32, 64, (char)128, 0, 0, 0, 0, 0, 0, 0, 0);
  const __m128i lut_hi = _mm_setr_epi8(0, 0, 0, 0, 0, 0, 0, 0, 1, 2, 4, 8, 16, 32, 64, (char)128);
  const __m128i nib = _mm_set1_epi8(0x0F);
  for (long long g = g0; g < g1; ++g) {
    const __m128i x = _mm_cvtsi64_si128((long long)boards[g]);
    const __m128i lo = _mm_and_si128(x, nib), hi = _mm_and_si128(_mm_srli_epi64(x, 4), nib);
    const __m128i e = _mm_unpacklo_epi8(lo, hi);                       // 16 exponents in cell order
    const __m128i vl = _mm_shuffle_epi8(lut_lo, e), vh = _mm_shuffle_epi8(lut_hi, e);
    const __m128i w0 = _mm_unpacklo_epi8(vl, vh), w1 = _mm_unpackhi_epi8(vl, vh);   // 8 x u16 each
    int32_t* o = obs + g * 16;
    _mm256_stream_si256(reinterpret_cast<__m256i*>(o), _mm256_cvtepu16_epi32(w0));
    _mm256_stream_si256(reinterpret_cast<__m256i*>(o + 8), _mm256_cvtepu16_epi32(w1));
  }
  _mm_sfence();
}

}  // namespace

void pulse_tfe_expand_range(const uint64_t* boards, int32_t* obs, long long g0, long long g1, int side) {
  if (side == 4 && have_avx2() && (reinterpret_cast<uintptr_t>(obs) & 31u) == 0) tfe_expand_avx2_side4(boards, obs, g0, g1);
  else tfe_expand_portable(boards, obs, g0, g1, side);
}
