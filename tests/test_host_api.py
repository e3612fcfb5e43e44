"""Host-side mirror of the reference's plugin API (games/*, utils, Node) against the
reference's golden outputs.  CPU only (no search)."""
import numpy as np
import pytest

import az_oracle as orc
from golden_util import load_json, board_from_words


def game_cls(name):
    from alpha_zero_b200 import games
    return {"tictactoe": games.TicTacToe, "connect4": games.Connect4, "gomoku": games.Gomoku}[name]


@pytest.mark.parametrize("name,g", [("rules_tictactoe.json", orc.TICTACTOE), ("rules_connect4.json", orc.CONNECT4),
                                    ("rules_gomoku7.json", orc.GOMOKU)])
def test_game_statics_follow_reference(name, g):
    G = game_cls(g.name)
    assert (G.rows, G.cols, G.action_dim, G.state_dim) == (g.rows, g.cols, g.action_dim, g.state_dim)
    for plies in load_json(name):
        board = G().board
        for rec in plies:
            assert np.array_equal(board[:2], board_from_words(g, rec["p0"], rec["p1"])[:2])
            legal = G.get_valid_moves(board)
            assert [r * g.cols + c for r, c in legal] == rec["legal"]
            mv = tuple(rec["move"])
            snap = board.copy()
            nxt = G.make_move(board, rec["player"], mv)
            assert nxt == 1 - rec["player"]
            if g.planes == 3:
                assert (board[2] == nxt).all()
            assert G.check_winner(board, rec["player"], mv) == rec["winner"]
            probe = board.copy()
            G.undo_move(probe, nxt, mv)
            assert np.array_equal(probe[:2], snap[:2])
            canon = G.get_canonical_board(board, nxt)
            assert np.array_equal(canon[0], board[nxt]) and np.array_equal(canon[1], board[1 - nxt])


def test_gomoku_resizes_like_the_reference_class():
    from alpha_zero_b200.games import Gomoku
    saved = (Gomoku.rows, Gomoku.cols, Gomoku.action_dim, Gomoku.state_dim)
    try:
        Gomoku.rows = Gomoku.cols = 15
        Gomoku.action_dim = Gomoku.state_dim = 225
        g = orc.gomoku_spec(15)
        for plies in load_json("rules_gomoku15.json"):
            for rec in plies[::7]:
                b = board_from_words(g, rec["p0"], rec["p1"])
                assert [r * 15 + c for r, c in Gomoku.get_valid_moves(b)] == rec["legal"]
    finally:
        Gomoku.rows, Gomoku.cols, Gomoku.action_dim, Gomoku.state_dim = saved


def test_node_policy_helpers():
    from alpha_zero_b200.ai import Node
    from alpha_zero_b200 import utils
    from alpha_zero_b200.games import Connect4
    root = Node(None, None, 0, 0)
    for col, n in ((0, 3), (2, 7), (5, 7)):
        ch = Node(root, (5, col), 1, 1, np.float32(0.1))
        ch.visit = n
        root.children.append(ch)
    root.visit = 18
    p = utils.get_probablity_distribution_of_children(root, Connect4)
    assert p.dtype == np.float64 and p.tolist() == [3 / 17, 0, 7 / 17, 0, 0, 7 / 17, 0]
    assert root.max_visit_child().prevAction == (5, 2)           # first maximum wins
    np.random.seed(0)
    assert root.sample_child(Connect4) in root.children
    root.refresh_ucb()
    assert "Visit: 7" in root.children[1].to_string(Connect4)
