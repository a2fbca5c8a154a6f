// Tetris 10x20 batched placement + line clear: per-env body of the step kernel.
// The reference has no Tetris behaviour (environments/Tetris/Tetris.py:1-13 and TetrisNumba.py:1-16
// are empty `pass` shells), so the rules are this repo's own spec, DESIGN.md section 6 (restated cell by
// cell in oracle/oracle_tetris.c, which is what the tests compare against):
//   board 10 x 20, row 0 at the bottom, rows[y] bit x = cell (x, y);
//   action a in 0..39: rotation a / 10, left edge a % 10 clamped to 10 - width; hard drop from above;
//   a piece that would rest with a cell at y >= 20 ends the game (board unchanged, reward 0);
//   otherwise lock, remove full rows, reward = rows removed, lines += reward, take the next piece.
// Here a piece/rotation is up to four 4-bit row masks, so collision with a board row is one shift and
// one AND, and a full row is `== 0x3FF`.
#pragma once
#include "common.cuh"

#define TET_W 10
#define TET_H 20
#define TET_FULL 0x3FFu
#define TET_SHAPES 28

// shape[piece * 4 + rot] = row masks (row k of the bounding box at bits 4k..4k+3) | width << 16 | height << 20.
// Pieces I O T S Z J L = 0..6; cells per rotation as listed in DESIGN.md section 6.
#define TET_S(mask, w, h) ((uint32_t)(mask) | ((uint32_t)(w) << 16) | ((uint32_t)(h) << 20))
#define TET_SHAPE_TABLE                                                                          \
  { TET_S(0x000F, 4, 1), TET_S(0x1111, 1, 4), TET_S(0x000F, 4, 1), TET_S(0x1111, 1, 4), /* I */  \
    TET_S(0x0033, 2, 2), TET_S(0x0033, 2, 2), TET_S(0x0033, 2, 2), TET_S(0x0033, 2, 2), /* O */  \
    TET_S(0x0027, 3, 2), TET_S(0x0131, 2, 3), TET_S(0x0072, 3, 2), TET_S(0x0232, 2, 3), /* T */  \
    TET_S(0x0063, 3, 2), TET_S(0x0132, 2, 3), TET_S(0x0063, 3, 2), TET_S(0x0132, 2, 3), /* S */  \
    TET_S(0x0036, 3, 2), TET_S(0x0231, 2, 3), TET_S(0x0036, 3, 2), TET_S(0x0231, 2, 3), /* Z */  \
    TET_S(0x0017, 3, 2), TET_S(0x0311, 2, 3), TET_S(0x0074, 3, 2), TET_S(0x0223, 2, 3), /* J */  \
    TET_S(0x0047, 3, 2), TET_S(0x0113, 2, 3), TET_S(0x0071, 3, 2), TET_S(0x0322, 2, 3)  /* L */ }

// does the shape, with its bounding box's bottom-left corner at (x, y), overlap a filled cell?
PULSE_HD bool tet_collides(const uint16_t* rows, uint32_t shape, int h, int x, int y) {
  bool hit = false;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int yy = y + k;
    if (k < h && yy < TET_H) hit = hit || ((uint32_t)rows[yy] & (((shape >> (4 * k)) & 15u) << x));
  }
  return hit;
}

// rows 2i and 2i+1 as one 32-bit word (the rows array is 4-byte aligned)
PULSE_HD uint32_t tet_row_pair(const uint16_t* rows, int i) {
#ifdef __CUDA_ARCH__
  return reinterpret_cast<const uint32_t*>(rows)[i];
#else
  return (uint32_t)rows[2 * i] | ((uint32_t)rows[2 * i + 1] << 16);
#endif
}

// One step of one env.  rows: this env's 20 rows (shared memory on the device).  A done env ignores the
// action (reward 0) unless auto_reset, in which case the step that finds it done clears the board,
// zeroes `lines` and takes `reset_piece`.  Returns true when rows changed.
PULSE_HD bool tet_step_one(uint16_t* rows, const uint32_t* shapes, int& piece, int32_t& lines, bool& done, long long action,
                           int next_piece, int reset_piece, int auto_reset, int32_t& reward) {
  reward = 0;
  if (done) {
    if (!auto_reset) return false;
#ifdef __CUDA_ARCH__
#pragma unroll
    for (int i = 0; i < TET_H / 2; ++i) reinterpret_cast<uint32_t*>(rows)[i] = 0u;
#else
    for (int y = 0; y < TET_H; ++y) rows[y] = 0;
#endif
    piece = reset_piece; lines = 0; done = false;
    return true;
  }
  int a = action < 0 ? 0 : action > 39 ? 39 : (int)action;
  const int rot = a / 10;
  const uint32_t shape = shapes[piece * 4 + rot];
  const int w = (shape >> 16) & 7, h = (shape >> 20) & 7;
  int x = a - rot * 10;
  if (x > TET_W - w) x = TET_W - w;
  // Hard drop.  Let `top` be the highest row with a filled cell in the piece's columns.  Dropping from
  // above, the piece is free at y = top + 1, and it cannot rest below top + 2 - h: the column that is
  // filled at `top` holds a piece cell at offset <= h - 1, which must end up above `top`.  So instead of
  // stepping down from row 20, find `top` (two rows per word) and test at most h - 1 positions.
  const uint32_t colmask = ((1u << w) - 1u) << x, colmask2 = colmask | (colmask << 16);
  int wi = TET_H / 2 - 1;
  uint32_t pair = 0;
  while (wi >= 0 && !((pair = tet_row_pair(rows, wi)) & colmask2)) --wi;
  const int top = wi < 0 ? -1 : ((pair >> 16) & colmask) ? 2 * wi + 1 : 2 * wi;
  int y = top + 1;
  const int ymin = top + 2 - h > 0 ? top + 2 - h : 0;
  while (y > ymin && !tet_collides(rows, shape, h, x, y - 1)) --y;
  if (y + h > TET_H) { done = true; return false; }     // locks out above the board: game over
  int cleared = 0;
#pragma unroll
  for (int k = 0; k < 4; ++k)
    if (k < h) {
      const uint32_t r = (uint32_t)rows[y + k] | (((shape >> (4 * k)) & 15u) << x);
      rows[y + k] = (uint16_t)r;
      cleared += r == TET_FULL;
    }
  if (cleared) {                                        // only rows y..y+h-1 can be full
    int wr = y;
    for (int r = y; r < TET_H; ++r) {
      const uint16_t v = rows[r];
      if (v != TET_FULL) rows[wr++] = v;
    }
    for (; wr < TET_H; ++wr) rows[wr] = 0;
  }
  reward = cleared; lines += cleared;
  piece = next_piece;
  return true;
}
