"""Gomoku plugin (reference: games/gomoku.py): n x n (7x7 in the reference), five in a row,
2 feature planes.  Legal moves are the empty 8-neighbours of existing stones (centre on an
empty board), returned -- like the reference -- in CPython `set` iteration order, which is the
child order of the search."""
import numpy as np

from . import game as _g
from .. import utils as _utils
from .selfplay import play_one_game


class Gomoku(_g.Game):
    rows, cols = 7, 7
    action_dim = rows * cols
    state_dim = rows * cols
    feature_dim = 2
    win_length = 5
    name = "gomoku"
    logger = _utils.get_game_logger("Gomoku")

    def __init__(self):
        self.board = np.zeros((Gomoku.feature_dim, Gomoku.rows, Gomoku.cols), dtype=np.float32)

    @staticmethod
    def display_board(board):
        _g.render(board, header=True)

    @staticmethod
    def get_canonical_board(board, current_player):
        return _g.swap_sides(board, current_player, keep_third_plane=False)

    @staticmethod
    def get_action_idx(action):
        return action[0] * Gomoku.rows + action[1]

    @staticmethod
    def make_move(board, current_player, action):
        r, c = action
        if board[0, r, c] != 0 or board[1, r, c] != 0:
            print("Invalid move. Try again.")
            return current_player
        board[current_player, r, c] = 1
        return 1 - current_player

    @staticmethod
    def undo_move(board, current_player, action):
        board[1 - current_player, action[0], action[1]] = 0

    @staticmethod
    def check_winner(board, player, action):
        return _g.line_win(board, player, action, 5)

    @staticmethod
    def get_valid_moves(board):
        rows, cols = Gomoku.rows, Gomoku.cols
        stones = (board[0] == 1) | (board[1] == 1)
        free = (board[0] == 0) & (board[1] == 0)
        found = set()                                   # a real set: its order IS the contract
        for r, c in zip(*np.nonzero(stones)):           # row-major scan of stones
            r, c = int(r), int(c)
            for dr, dc in ((0, 1), (1, 0), (1, 1), (1, -1)):
                for sgn in (1, -1):
                    rr, cc = r + sgn * dr, c + sgn * dc
                    if 0 <= rr < rows and 0 <= cc < cols and free[rr, cc]:
                        found.add((rr, cc))
        if not found:
            found.add((rows // 2, cols // 2))
        return list(found)

    @staticmethod
    def mcts(model, board, root, mcts_iterations, dirichlet=True):
        from ..ai import MCTS
        MCTS.mcts(model, board, root, Gomoku, mcts_iterations, dirichlet)

    @staticmethod
    def get_input(board):
        r, c = map(int, input("Enter row and column: ").split())
        return None if board[0, r, c] == 1 or board[1, r, c] == 1 else (r, c)

    def self_play(self, model, mcts_iter, display=False):
        return play_one_game(Gomoku, self.board, model, mcts_iter, display, sample_until=8)   # gomoku.py:143
