"""Host-side board object with the reference's public surface (gomoku_board.py:17-327 there).

This is the *input type* of the hot path, not the hot path: the engine converts it to two 256-bit
bitboards at the boundary (engine.Engine.boards_from_cells) and all rule logic that matters for
throughput runs in the CUDA kernels.  The class keeps the attributes callers read and write
(`board`, `move_history`, `current_player`, `game_over`, `winner`) and the same method names and
return conventions, so `training.play_one_game` / `game_ui` style code runs unchanged.
"""
import numpy as np


class GomokuBoard:
    BOARD_SIZE = 15
    EMPTY, BLACK, WHITE = 0, 1, 2
    DIRECTIONS = [(0, 1), (1, 0), (1, 1), (1, -1)]
    MAX_MOVES = 200          # the reference ends a game as a draw at 200 stones (gomoku_board.py:106)

    def __init__(self):
        n = self.BOARD_SIZE
        self.board = np.zeros((n, n), dtype=int)
        self.move_history = []
        self.current_player = self.BLACK
        self.game_over = False
        self.winner = None

    # ---- state management
    def reset(self):
        self.board[:] = self.EMPTY
        del self.move_history[:]
        self.current_player = self.BLACK
        self.game_over = False
        self.winner = None

    def copy_board(self):
        other = GomokuBoard.__new__(GomokuBoard)
        other.board = self.board.copy()
        other.move_history = list(self.move_history)
        other.current_player = self.current_player
        other.game_over = self.game_over
        other.winner = self.winner
        return other

    # ---- rules
    def _inside(self, row, col):
        return 0 <= row < self.BOARD_SIZE and 0 <= col < self.BOARD_SIZE

    def is_valid_move(self, row, col):
        return self._inside(row, col) and self.board[row, col] == self.EMPTY

    def _count_consecutive(self, row, col, direction, player):
        dr, dc = direction
        k = 0
        r, c = row + dr, col + dc
        while self._inside(r, c) and self.board[r, c] == player:
            k += 1
            r += dr
            c += dc
        return k

    def check_win(self, row, col):
        """>= 5 contiguous stones through (row, col); overlines count."""
        who = self.board[row, col]
        for dr, dc in self.DIRECTIONS:
            run = 1 + self._count_consecutive(row, col, (dr, dc), who) \
                    + self._count_consecutive(row, col, (-dr, -dc), who)
            if run >= 5:
                return True
        return False

    def is_board_full(self):
        return not (self.board == self.EMPTY).any()

    def make_move(self, row, col):
        if self.game_over or not self.is_valid_move(row, col):
            return False
        mover = self.current_player
        self.board[row, col] = mover
        self.move_history.append((row, col, mover))
        if self.check_win(row, col):
            self.game_over, self.winner = True, mover
        elif self.is_board_full() or len(self.move_history) >= self.MAX_MOVES:
            self.game_over, self.winner = True, None
        self.current_player = self.BLACK + self.WHITE - mover
        return True

    def undo_move(self):
        if not self.move_history:
            return False
        row, col, mover = self.move_history.pop()
        self.board[row, col] = self.EMPTY
        self.game_over, self.winner = False, None
        self.current_player = mover
        return True

    # ---- queries
    def get_valid_moves(self):
        """Every empty cell in raster order, whether or not the game is over."""
        rows, cols = np.nonzero(self.board == self.EMPTY)
        return list(zip(rows.tolist(), cols.tolist()))

    def get_board_state(self):
        return self.board.copy()

    def get_board_tensor(self):
        """(3, 15, 15) float32 planes: black, white, empty (absolute colours)."""
        return np.stack([self.board == self.BLACK, self.board == self.WHITE,
                         self.board == self.EMPTY]).astype(np.float32)

    def get_last_move(self):
        return self.move_history[-1][:2] if self.move_history else None

    def get_move_count(self):
        return len(self.move_history)

    def get_game_status(self):
        return {"current_player": self.current_player, "game_over": self.game_over,
                "winner": self.winner, "move_count": self.get_move_count(),
                "board_full": self.is_board_full()}

    # ---- engine boundary
    def to_cells(self):
        """uint8[225] raster cells, the layout gz_boards_from_cells takes."""
        return np.ascontiguousarray(self.board, dtype=np.uint8).reshape(225)

    def __str__(self):
        glyph = {self.EMPTY: ".", self.BLACK: "●", self.WHITE: "○"}
        head = "   " + " ".join("%2d" % i for i in range(self.BOARD_SIZE))
        body = ["%2d " % i + " ".join(glyph[int(v)] for v in row) for i, row in enumerate(self.board)]
        return "\n".join([head] + body)

    def __eq__(self, other):
        return (isinstance(other, GomokuBoard) and self.current_player == other.current_player
                and np.array_equal(self.board, other.board))

    __hash__ = None
