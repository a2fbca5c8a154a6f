"""TEST INFRASTRUCTURE ONLY.  A second, independently written implementation of the Tetris rules of DESIGN.md section 6,
in plain Python on a list-of-lists grid.  It shares no code, table or representation with oracle/oracle_tetris.c
(bit rows, packed shape words) or pulselib_b200/csrc/tetris_body.cuh: the piece shapes below were typed from the table
in DESIGN.md, the drop is a literal "move down while it fits" loop.  The reference ships only empty Tetris shells
(/root/reference/environments/Tetris/Tetris.py:1-13), so parity with the reference stays UNPINNED for this env; what
this file buys is that two implementations written by different routes agree on the spec."""

W, H = 10, 20

# cells (x, y) anchored at the bounding box's bottom-left corner, pieces I O T S Z J L = 0..6
I0 = [(0, 0), (1, 0), (2, 0), (3, 0)]
I1 = [(0, 0), (0, 1), (0, 2), (0, 3)]
O0 = [(0, 0), (1, 0), (0, 1), (1, 1)]
S0 = [(0, 0), (1, 0), (1, 1), (2, 1)]
S1 = [(1, 0), (1, 1), (0, 1), (0, 2)]
Z0 = [(1, 0), (2, 0), (0, 1), (1, 1)]
Z1 = [(0, 0), (0, 1), (1, 1), (1, 2)]
SHAPES = [
    [I0, I1, I0, I1],
    [O0, O0, O0, O0],
    [[(0, 0), (1, 0), (2, 0), (1, 1)], [(0, 0), (0, 1), (0, 2), (1, 1)], [(0, 1), (1, 1), (2, 1), (1, 0)], [(1, 0), (1, 1), (1, 2), (0, 1)]],
    [S0, S1, S0, S1],
    [Z0, Z1, Z0, Z1],
    [[(0, 0), (1, 0), (2, 0), (0, 1)], [(0, 0), (0, 1), (0, 2), (1, 2)], [(2, 0), (0, 1), (1, 1), (2, 1)], [(0, 0), (1, 0), (1, 1), (1, 2)]],
    [[(0, 0), (1, 0), (2, 0), (2, 1)], [(0, 0), (1, 0), (0, 1), (0, 2)], [(0, 0), (0, 1), (1, 1), (2, 1)], [(1, 0), (1, 1), (1, 2), (0, 2)]],
]


class TetrisSpec:
    """One env.  grid[y][x], row 0 at the bottom."""

    def __init__(self, first_piece):
        self.grid = [[False] * W for _ in range(H)]
        self.piece, self.lines, self.done = int(first_piece), 0, False

    def rows_as_bits(self):
        return [sum(1 << x for x in range(W) if self.grid[y][x]) for y in range(H)]

    def _fits(self, cells, x0, y0):
        for cx, cy in cells:
            x, y = x0 + cx, y0 + cy
            if y < 0:
                return False
            if y < H and self.grid[y][x]:
                return False
        return True

    def step(self, action, next_piece, reset_piece, auto_reset):
        """-> reward.  Rules: DESIGN.md section 6."""
        if self.done:
            if auto_reset:
                self.grid = [[False] * W for _ in range(H)]
                self.lines, self.piece, self.done = 0, int(reset_piece), False
            return 0
        a = min(max(int(action), 0), 39)
        cells = SHAPES[self.piece][a // 10]
        width = 1 + max(cx for cx, _ in cells)
        x0 = min(a % 10, W - width)
        y0 = H + 4                      # well above the stack: drop straight down while the next position fits
        while self._fits(cells, x0, y0 - 1):
            y0 -= 1
        if any(y0 + cy >= H for _, cy in cells):
            self.done = True            # a cell would rest above the board: lock-out, board unchanged
            return 0
        for cx, cy in cells:
            self.grid[y0 + cy][x0 + cx] = True
        kept = [row for row in self.grid if not all(row)]
        cleared = H - len(kept)
        self.grid = kept + [[False] * W for _ in range(cleared)]
        self.lines += cleared
        self.piece = int(next_piece)
        return cleared
