"""Connect4 plugin (reference: games/connect4.py): 6x7, four in a row, actions are
(drop_row, col) tuples indexed by column, 3 feature planes (plane 2 = side to move)."""
import numpy as np

from . import game as _g
from .. import utils as _utils
from .selfplay import play_one_game


class Connect4(_g.Game):
    rows, cols = 6, 7
    action_dim = 7
    state_dim = 42
    feature_dim = 3
    win_length = 4
    name = "connect4"
    logger = _utils.get_game_logger("connect4")

    def __init__(self):
        self.board = np.zeros((3, 6, 7), dtype=np.float32)

    @staticmethod
    def display_board(board):
        _g.render(board, header=True)

    @staticmethod
    def get_action_idx(move):
        return move[1]

    @staticmethod
    def get_drop_row(board, col):
        """Lowest empty row of a column (row 5 is the bottom), None when full."""
        free = np.nonzero((board[0, :, col] == 0) & (board[1, :, col] == 0))[0]
        return int(free[-1]) if len(free) else None

    @staticmethod
    def get_valid_moves(board):
        moves = []
        for col in range(7):                            # columns ascending = child order
            if board[0, 0, col] == 0 and board[1, 0, col] == 0:
                moves.append((Connect4.get_drop_row(board, col), col))
        return moves

    @staticmethod
    def make_move(board, current_player, action):
        r, c = action
        if r is None:
            print("Invalid move: column is full.")
            return current_player
        board[current_player, r, c] = 1
        board[2] = 1 - current_player
        return 1 - current_player

    @staticmethod
    def undo_move(board, current_player, action):
        board[1 - current_player, action[0], action[1]] = 0
        board[2] = 1 - current_player

    @staticmethod
    def check_winner(board, player, action):
        return _g.line_win(board, player, action, 4)

    @staticmethod
    def get_canonical_board(board, current_player):
        return _g.swap_sides(board, current_player, keep_third_plane=True)

    @staticmethod
    def mcts(model, board, root, mcts_iterations, dirichlet=True):
        from ..ai import MCTS
        MCTS.mcts(model, board, root, Connect4, mcts_iterations, dirichlet)

    @staticmethod
    def get_input(board):
        col = int(input("Enter column (0~6): "))
        row = Connect4.get_drop_row(board, col)
        return None if row is None else (row, col)

    def self_play(self, model, mcts_iter, display=False):
        boards, _actions, policies, _qs, winner = play_one_game(Connect4, self.board, model, mcts_iter, display,
                                                                sample_until=None)
        return boards, policies, winner                 # 3-tuple, games/connect4.py:151
