"""TicTacToe plugin (reference: games/tictactoe.py): 3x3, three in a row, 3 feature planes
(plane 2 = id of the side to move)."""
import numpy as np

from . import game as _g
from .. import utils as _utils
from .selfplay import play_one_game


class TicTacToe(_g.Game):
    rows, cols = 3, 3
    action_dim = 9
    state_dim = 9
    feature_dim = 3          # the reference omits this attribute; main.py:134 needs it
    win_length = 3
    name = "tictactoe"
    logger = _utils.get_game_logger("tictactoe")

    def __init__(self):
        self.board = np.zeros((3, 3, 3), dtype=np.float32)

    @staticmethod
    def display_board(board):
        _g.render(board, header=False)

    @staticmethod
    def get_action_idx(action):
        return action[0] * 3 + action[1]

    @staticmethod
    def get_valid_moves(board):
        free = (board[0] == 0) & (board[1] == 0)
        return [(int(r), int(c)) for r, c in zip(*np.nonzero(free))]        # row-major

    @staticmethod
    def make_move(board, current_player, action):
        r, c = action
        if board[0, r, c] != 0 or board[1, r, c] != 0:
            print("Invalid move. Try again.")
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
        return _g.line_win(board, player, action, 3)

    @staticmethod
    def get_canonical_board(board, current_player):
        return _g.swap_sides(board, current_player, keep_third_plane=True)

    @staticmethod
    def mcts(model, board, root, mcts_iterations, dirichlet=True):
        from ..ai import MCTS
        MCTS.mcts(model, board, root, TicTacToe, mcts_iterations, dirichlet)

    @staticmethod
    def get_input(board):
        r, c = map(int, input("Enter row and column: ").split())
        return None if board[0, r, c] == 1 or board[1, r, c] == 1 else (r, c)

    def self_play(self, model, mcts_iter, display=False):
        boards, _actions, policies, _qs, winner = play_one_game(TicTacToe, self.board, model, mcts_iter, display,
                                                                sample_until=None)
        return boards, policies, winner                 # 3-tuple, games/tictactoe.py:133
