from .game import Game  # noqa: F401
from .tictactoe import TicTacToe  # noqa: F401
from .connect4 import Connect4  # noqa: F401
from .gomoku import Gomoku  # noqa: F401
