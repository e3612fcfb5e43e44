"""Game plugin interface of the reference (games/game.py:4-38) plus the shared host-side
rule code of the three games.

The reference passes a game CLASS around as a namespace of static rule functions on
caller-owned float32 boards [C, H, W] that are mutated in place.  That contract is kept.
These statics serve the host driver (one call per ply); the search itself never calls
them -- inside MCTS the rules are bitboard kernels (csrc/az_rules.cuh).
"""
from abc import ABC, abstractmethod

import numpy as np

_LINES = ((0, 1), (1, 0), (1, 1), (1, -1))


class Game(ABC):
    """Abstract statics every game provides (same names and argument order as the reference)."""

    @staticmethod
    @abstractmethod
    def display_board(board):
        ...

    @staticmethod
    @abstractmethod
    def get_action_idx(action):
        ...

    @staticmethod
    @abstractmethod
    def get_valid_moves(board):
        ...

    @staticmethod
    @abstractmethod
    def make_move(board, current_player, action):
        ...

    @staticmethod
    @abstractmethod
    def undo_move(board, current_player, action):
        ...

    @staticmethod
    @abstractmethod
    def check_winner(board, player, action):
        ...


def line_win(board, player, action, need):
    """`player` if a run of >= need of its stones passes through `action`, else -1
    (games/gomoku.py:65-91 and siblings; overlines count)."""
    _, rows, cols = board.shape
    plane = board[player]
    r0, c0 = action
    for dr, dc in _LINES:
        run = 1
        for sgn in (1, -1):
            r, c = r0 + sgn * dr, c0 + sgn * dc
            while 0 <= r < rows and 0 <= c < cols and plane[r, c] == 1:
                run += 1
                r += sgn * dr
                c += sgn * dc
        if run >= need:
            return player
    return -1


def swap_sides(board, current_player, keep_third_plane):
    """Canonical form: side to move in plane 0 (games/gomoku.py:34-40; for the 3-plane games
    the author's legacy version, ai/mcts.py:126-137: swap 0/1, keep plane 2)."""
    if current_player == 0:
        return board
    out = np.empty_like(board)
    out[0], out[1] = board[1], board[0]
    if keep_third_plane:
        out[2] = board[2]
    return out


def render(board, header):
    rows, cols = board.shape[1:]
    grid = np.full((rows, cols), " ")
    grid[board[0] == 1] = "O"
    grid[board[1] == 1] = "X"
    if header:
        print("\n  " + " ".join(str(c % 10) for c in range(cols)))
    for i, row in enumerate(grid):
        print((str(i % 10) + " " if header else "") + " ".join(row))
    print()
