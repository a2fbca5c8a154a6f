/* TEST INFRASTRUCTURE ONLY -- see oracle_common.h.
 * Tetris has NO behaviour in the reference (environments/Tetris/Tetris.py:1-13 and
 * TetrisNumba.py:1-16 are empty `pass` shells), so there is nothing to restate: this file IS the
 * specification of the batched 10x20 placement + line-clear env named by BASELINE.json configs[3]
 * (DESIGN.md section 6), written cell by cell from (x, y) cell lists (no row-mask piece tables, no
 * column-height shortcuts) so that it shares no logic with the bit-row CUDA kernel.  Parity with the reference: unpinned (no reference semantics exist).
 *
 * Spec.  Board: 10 columns x 20 rows, row 0 at the bottom.  Seven tetrominoes (I O T S Z J L = 0..6),
 * four rotation states each, given below as four (x, y) cells anchored at the bounding box's
 * bottom-left corner.  Action a in 0..39: rotation = a / 10, column = a % 10 = x of the bounding
 * box's left edge, clamped to 10 - width so every action is placeable.  The piece is dropped straight
 * down from above the board and rests at the lowest y it reaches without overlapping a filled cell.
 * If any of its cells would rest at y >= 20 the game is over: done = 1, board unchanged, reward 0.
 * Otherwise the cells are filled, full rows are removed (rows above move down), reward = number of
 * rows removed (0..4), lines += reward, and the next piece is taken from the chance input.  A done
 * env ignores actions (reward 0) unless auto_reset is set, in which case the step that finds it done
 * clears the board, zeroes `lines`, takes a fresh piece and returns reward 0, done 0. */
#include "oracle_common.h"

#define TW 10
#define TH 20

/* cells[piece][rot][k] = {x, y} */
static const int8_t CELLS[7][4][4][2] = {
  /* I */ {{{0,0},{1,0},{2,0},{3,0}}, {{0,0},{0,1},{0,2},{0,3}}, {{0,0},{1,0},{2,0},{3,0}}, {{0,0},{0,1},{0,2},{0,3}}},
  /* O */ {{{0,0},{1,0},{0,1},{1,1}}, {{0,0},{1,0},{0,1},{1,1}}, {{0,0},{1,0},{0,1},{1,1}}, {{0,0},{1,0},{0,1},{1,1}}},
  /* T */ {{{0,0},{1,0},{2,0},{1,1}}, {{0,0},{0,1},{0,2},{1,1}}, {{0,1},{1,1},{2,1},{1,0}}, {{1,0},{1,1},{1,2},{0,1}}},
  /* S */ {{{0,0},{1,0},{1,1},{2,1}}, {{1,0},{1,1},{0,1},{0,2}}, {{0,0},{1,0},{1,1},{2,1}}, {{1,0},{1,1},{0,1},{0,2}}},
  /* Z */ {{{1,0},{2,0},{0,1},{1,1}}, {{0,0},{0,1},{1,1},{1,2}}, {{1,0},{2,0},{0,1},{1,1}}, {{0,0},{0,1},{1,1},{1,2}}},
  /* J */ {{{0,0},{1,0},{2,0},{0,1}}, {{0,0},{0,1},{0,2},{1,2}}, {{2,0},{0,1},{1,1},{2,1}}, {{0,0},{1,0},{1,1},{1,2}}},
  /* L */ {{{0,0},{1,0},{2,0},{2,1}}, {{0,0},{1,0},{0,1},{0,2}}, {{0,0},{0,1},{1,1},{2,1}}, {{1,0},{1,1},{1,2},{0,2}}},
};

static int filled(const uint16_t* rows, int x, int y) { return y < TH && ((rows[y] >> x) & 1u); }

static int collides(const uint16_t* rows, int piece, int rot, int x, int y) {
  for (int k = 0; k < 4; ++k)
    if (filled(rows, x + CELLS[piece][rot][k][0], y + CELLS[piece][rot][k][1])) return 1;
  return 0;
}

/* next piece of env `env_id` at global step counter `t`: word 0 of the block; a reset taken at
 * counter t uses word 1 for its first piece */
void oracle_tetris_tape_pieces(uint64_t seed, uint64_t env_base, uint32_t counter, int word, int n, uint8_t* out) {
  for (int g = 0; g < n; ++g) {
    uint32_t blk[4];
    oracle_tape(seed, env_base + (uint64_t)g, ORACLE_STREAM_TETRIS_PIECE, counter, blk);
    out[g] = (uint8_t)oracle_mulhi32(blk[word], 7u);
  }
}

/* rows: uint16 [N,20] (bit x of rows[y] = cell (x,y)); piece: uint8 [N] current piece */
void oracle_tetris_reset(uint16_t* rows, uint8_t* piece, int32_t* lines, uint8_t* done, const uint8_t* first_piece, int n) {
  for (int g = 0; g < n; ++g) {
    for (int y = 0; y < TH; ++y) rows[(size_t)g * TH + y] = 0;
    piece[g] = first_piece[g]; lines[g] = 0; done[g] = 0;
  }
}

void oracle_tetris_step(uint16_t* rows_all, uint8_t* piece, int32_t* lines, uint8_t* done, const int64_t* actions,
                        const uint8_t* next_piece, const uint8_t* reset_piece, int32_t* rewards, int auto_reset, int n) {
#pragma omp parallel for schedule(static) if (n > 16384)
  for (int g = 0; g < n; ++g) {
    uint16_t* rows = rows_all + (size_t)g * TH;
    rewards[g] = 0;
    if (done[g]) {
      if (auto_reset) {
        for (int y = 0; y < TH; ++y) rows[y] = 0;
        piece[g] = reset_piece[g]; lines[g] = 0; done[g] = 0;
      }
      continue;
    }
    int a = (int)actions[g];
    if (a < 0) a = 0;
    if (a > 39) a = 39;
    const int p = piece[g], rot = a / 10;
    int w = 0, h = 0;
    for (int k = 0; k < 4; ++k) {
      if (CELLS[p][rot][k][0] + 1 > w) w = CELLS[p][rot][k][0] + 1;
      if (CELLS[p][rot][k][1] + 1 > h) h = CELLS[p][rot][k][1] + 1;
    }
    int x = a % 10;
    if (x > TW - w) x = TW - w;
    int y = TH;                                   /* entirely above the board */
    while (y > 0 && !collides(rows, p, rot, x, y - 1)) --y;
    if (y + h > TH) { done[g] = 1; continue; }    /* locks out above the board: game over */
    for (int k = 0; k < 4; ++k) rows[y + CELLS[p][rot][k][1]] |= (uint16_t)(1u << (x + CELLS[p][rot][k][0]));
    int wr = 0, cleared = 0;
    for (int r = 0; r < TH; ++r) {
      if (rows[r] == 0x3FF) { ++cleared; continue; }
      rows[wr++] = rows[r];
    }
    for (; wr < TH; ++wr) rows[wr] = 0;
    rewards[g] = cleared; lines[g] += cleared;
    piece[g] = next_piece[g];
  }
}
