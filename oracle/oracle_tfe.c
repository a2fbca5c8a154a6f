/* TEST INFRASTRUCTURE ONLY -- see oracle_common.h.
 * CPU restatement of /root/reference/environments/2048/TFE.py on plain int32 boards of TILE VALUES
 * (2, 4, 8 ...), exactly the reference's representation (:121): numba_rotate_with_buffer :38-44,
 * squash_row_optimized :75-108, add_tile_numba :17-34, is_game_over_numba :48-67, TFE.reset
 * :143-149, TFE.step :152-189.  The reference steps ONE board per object; this applies the same
 * per-board rules to a batch.  The only change is the source of chance: add_tile's two numba RNG
 * draws (random.randint(0, n_empty-1), random.random() > 0.9) are replaced by two 32-bit words
 * (idx = mulhi32(u0, n_empty), four = u1 > TFE_FOUR_THRESHOLD) so that the reference, this oracle and
 * the CUDA kernel can be fed the same draws (tests/golden/make_golden.py patches add_tile_numba in
 * the loaded reference module with the same rule).  Square boards, side 2..4. */
#include "oracle_common.h"

#define TFE_FOUR_THRESHOLD 3865470566u /* floor(0.9 * 2^32): u1 > this <=> "random() > 0.9" */
#define TFE_MAX_SIDE 4

/* numba_rotate_with_buffer (:38-44): out[m-1-j, i] = board[i, j]  (one counter-clockwise quarter turn) */
static void rotate(const int32_t* b, int32_t* out, int n) {
  for (int i = 0; i < n; ++i)
    for (int j = 0; j < n; ++j) out[(n - 1 - j) * n + i] = b[i * n + j];
}

/* squash_row_optimized (:75-108); returns the row's score */
static int32_t squash_row(const int32_t* row, int32_t* result, int n, int* changed) {
  int32_t score = 0;
  for (int i = 0; i < n; ++i) result[i] = 0;
  int w = 0, last_merged = 0;
  for (int i = 0; i < n; ++i) {
    int32_t v = row[i];
    if (v != 0) {
      if (result[w] == 0) result[w] = v;
      else if (result[w] == v && !last_merged) { result[w] = v * 2; score += v * 2; last_merged = 1; }
      else { ++w; result[w] = v; last_merged = 0; }
    }
  }
  *changed = 0;
  for (int i = 0; i < n; ++i) if (row[i] != result[i]) { *changed = 1; break; }
  return score;
}

/* add_tile_numba (:17-34) with injected draws */
static void add_tile(int32_t* b, int n, uint32_t u0, uint32_t u1) {
  int cells[TFE_MAX_SIDE * TFE_MAX_SIDE], ne = 0;
  for (int i = 0; i < n * n; ++i) if (b[i] == 0) cells[ne++] = i;  /* row-major order */
  if (ne == 0) return;
  int idx = (int)oracle_mulhi32(u0, (uint32_t)ne);
  b[cells[idx]] = u1 > TFE_FOUR_THRESHOLD ? 4 : 2;
}

/* is_game_over_numba (:48-67) */
static int game_over(const int32_t* b, int n) {
  for (int i = 0; i < n * n; ++i) if (b[i] == 0) return 0;
  for (int i = 0; i < n; ++i) for (int j = 0; j + 1 < n; ++j) if (b[i * n + j] == b[i * n + j + 1]) return 0;
  for (int i = 0; i + 1 < n; ++i) for (int j = 0; j < n; ++j) if (b[i * n + j] == b[(i + 1) * n + j]) return 0;
  return 1;
}

static int bit_length(int64_t v) { int n = 0; while (v > 0) { ++n; v >>= 1; } return n; }

/* chance words of the tape for a batch: words_per_env of the 4-word block at `counter` */
void oracle_tfe_tape_chance(uint64_t seed, uint64_t env_base, uint32_t stream, uint32_t counter, int n,
                            int words_per_env, uint32_t* out) {
  for (int g = 0; g < n; ++g) {
    uint32_t blk[4];
    oracle_tape(seed, env_base + (uint64_t)g, stream, counter, blk);
    for (int k = 0; k < words_per_env; ++k) out[(size_t)g * words_per_env + k] = blk[k];
  }
}

/* TFE.reset (:143-149): zero board, score 0, two spawns.  chance: [N,4] = (u0,u1) of spawn 1, (u0,u1) of spawn 2 */
void oracle_tfe_reset(int32_t* boards, const uint32_t* chance, int64_t* scores, uint8_t* done, int n_envs, int side) {
  for (int g = 0; g < n_envs; ++g) {
    int32_t* b = boards + (size_t)g * side * side;
    for (int i = 0; i < side * side; ++i) b[i] = 0;
    scores[g] = 0; done[g] = 0;
    add_tile(b, side, chance[4 * g], chance[4 * g + 1]);
    add_tile(b, side, chance[4 * g + 2], chance[4 * g + 3]);
  }
}

/* TFE.step (:152-189).  chance: [N,2] spawn draws.  auto_reset (batched-engine extension, not in the
 * reference): a board that was game-over when the step began is reset instead, using
 * reset_chance [N,4]; that step returns reward 0, done 0. */
void oracle_tfe_step(int32_t* boards, const int64_t* actions, const uint32_t* chance, int64_t* scores,
                     int32_t* rewards, uint8_t* done, int n_envs, int side, int auto_reset,
                     const uint32_t* reset_chance) {
  const int n = side;
#pragma omp parallel for schedule(static) if (n_envs > 16384)
  for (int g = 0; g < n_envs; ++g) {
    int32_t* board = boards + (size_t)g * n * n;
    if (auto_reset && done[g]) {
      for (int i = 0; i < n * n; ++i) board[i] = 0;
      scores[g] = 0; done[g] = 0; rewards[g] = 0;
      add_tile(board, n, reset_chance[4 * g], reset_chance[4 * g + 1]);
      add_tile(board, n, reset_chance[4 * g + 2], reset_chance[4 * g + 3]);
      continue;
    }
    int32_t a[16], t[16], s[16];
    for (int i = 0; i < n * n; ++i) a[i] = board[i];
    int k = (int)(actions[g] & 3);                      /* rotations_needed (:125) */
    for (int r = 0; r < k; ++r) { rotate(a, t, n); for (int i = 0; i < n * n; ++i) a[i] = t[i]; }  /* :158-163 */
    int64_t step_score = 0;
    for (int r = 0; r < n; ++r) { int ch; step_score += squash_row(a + r * n, s + r * n, n, &ch); }  /* :166-167 */
    scores[g] += step_score;                            /* :168 */
    for (int r = 0; r < (4 - k) % 4; ++r) { rotate(s, t, n); for (int i = 0; i < n * n; ++i) s[i] = t[i]; }  /* :171-178 */
    for (int i = 0; i < n * n; ++i) board[i] = s[i];    /* :181 */
    add_tile(board, n, chance[2 * g], chance[2 * g + 1]);  /* :182 always, even if nothing moved */
    rewards[g] = step_score > 0 ? bit_length(step_score) - 1 : 0;  /* :185-187 */
    done[g] = (uint8_t)game_over(board, n);             /* :189 */
  }
}
