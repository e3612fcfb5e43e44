"""Vanilla MCTS on the GPU (MCTS.mcts(None, ...): UCB1 + random rollouts, ai/mcts.py:57-79, utils.py:36,43).

Selection / expansion / backup are deterministic given the rollout results, so with the hash value in place of the
rollout (the reference's MCTS.simulate replaced the same way by oracle/gen_golden.py) visit counts and W must be
bit-exact -- ln(N) comes from the host C library exactly like math.log.  The rollouts themselves draw from another
RNG than NumPy's legacy generator, so they are checked statistically (tolerances in the tests)."""
import numpy as np
import pytest

import az_oracle as orc
import c_oracle as co
from golden_util import load_json, spec_of, position_after, fhex
from gpu_util import words_from_cells, random_positions, compare_forest_with_oracle

pytestmark = pytest.mark.gpu


def az():
    import alpha_zero_b200
    return alpha_zero_b200


def test_ucb1_search_matches_reference_goldens():
    for case in load_json("search_vanilla.json"):
        g = spec_of(case)
        board, player, mc = position_after(g, case["opening"])
        f = az().Forest(az().game_spec(g.name, g.rows), 1, case["iters"], vl=1)
        f.set_roots_words(words_from_cells(g, co.cells_from_planes(board)[None]), [player], [mc])
        f.set_mode(True)
        f.search_hash(case["iters"], case["seed"], case["qbits"])
        ch, ref = f.root_children(0), case["result"]
        ai = ch["cells"] % g.cols if g.gid == 1 else ch["cells"]
        assert ai.tolist() == ref["actions"]
        assert ch["visits"].tolist() == ref["visits"]
        assert ch["W"].tolist() == [fhex(v) for v in ref["values"]]
        assert ch["root_visit"] == ref["root_visit"] and ch["root_value"] == fhex(ref["root_value"])
        f.close()


@pytest.mark.parametrize("g,n,iters,vl", [(orc.CONNECT4, 512, 400, 1), (orc.GOMOKU, 256, 300, 1), (orc.TICTACTOE, 256, 200, 1),
                                          (orc.CONNECT4, 128, 300, 4)])
def test_ucb1_forest_matches_c_oracle(g, n, iters, vl):
    rng = np.random.default_rng(3)
    cells, players, mcs = random_positions(g, n, rng, 8 if g.gid else 4)
    co.set_ucb1(True)
    try:
        ref = co.search_batch(g, cells, players, mcs, iters, vl, seed=9, qbits=16)
    finally:
        co.set_ucb1(False)
    f = az().Forest(az().game_spec(g.name, g.rows), n, iters, vl=vl)
    f.set_roots_words(words_from_cells(g, cells), players, mcs)
    f.set_mode(True)
    f.search_hash(iters, 9, 16)
    compare_forest_with_oracle(f, ref)
    f.set_mode(False)                                           # and the forest goes back to PUCT
    ref2 = co.search_batch(g, cells, players, mcs, iters, vl, seed=9, qbits=16)
    f.search_hash(iters, 9, 16)
    compare_forest_with_oracle(f, ref2)
    f.close()


def rollout_values(g, cells, players, mcs, seed):
    n = cells.shape[0]
    f = az().Forest(az().game_spec(g.name, g.rows), n, 4, vl=1)
    f.set_roots_words(words_from_cells(g, cells), players, mcs)
    f.set_mode(True)
    f.begin_move(4)
    f.advance()                                                  # every root is a pending leaf now
    f.eval_rollout(seed, 0)
    r = f.read_leaves()
    assert r["n"] == n
    vals = r["values"].copy()
    assert (r["priors"] == 0).all()                              # node.py:58: children without priors
    f.close()
    return vals


def test_rollout_statistics_tictactoe():
    """Uniformly random play from the empty board: the side to move wins 58.5 %, draws 12.7 %, loses 28.8 %."""
    g, n = orc.TICTACTOE, 16384
    vals = rollout_values(g, np.zeros((n, 9), np.uint8), np.zeros(n, np.int32), np.zeros(n, np.int32), 1234)
    assert set(np.unique(vals).tolist()) <= {-1.0, 0.0, 1.0}
    win, draw, lose = (vals == 1).mean(), (vals == 0).mean(), (vals == -1).mean()
    assert abs(win - 0.585) < 0.02 and abs(draw - 0.127) < 0.02 and abs(lose - 0.288) < 0.02, (win, draw, lose)


@pytest.mark.parametrize("g,plies", [(orc.CONNECT4, 6), (orc.GOMOKU, 10)])
def test_rollout_statistics_match_the_oracle_rollout(g, plies):
    """Same positions, many rollouts each side: mean result of the device rollouts vs oracle.rollout
    (ai/mcts.py:62-79 restated) within 0.06 (about 4 sigma of the two 3000-sample means)."""
    rng = np.random.default_rng(5)
    cells1, players1, mcs1 = random_positions(g, 1, rng, plies)
    reps = 3000
    vals = rollout_values(g, np.repeat(cells1, reps, 0), np.repeat(players1, reps), np.repeat(mcs1, reps), 77)
    board = np.zeros((g.planes, g.rows, g.cols), np.float32)
    board[0].reshape(-1)[:] = cells1[0] == 1
    board[1].reshape(-1)[:] = cells1[0] == 2
    if g.planes == 3:
        board[2] = players1[0]
    node = orc.Node(None, None, int(players1[0]), int(mcs1[0]))
    r = np.random.RandomState(1)
    # oracle result: +1 if the player who moved INTO the node wins = -1 for the side to move
    ora = np.array([-orc.rollout(g, board, node, r) for _ in range(reps)], np.float64)
    assert abs(vals.mean() - ora.mean()) < 0.06, (vals.mean(), ora.mean())
    assert abs((vals == 0).mean() - (ora == 0).mean()) < 0.04


def test_mcts_api_vanilla_mode():
    """MCTS.mcts(None, ...) through the reference-facing API: finds the immediate win, leaves the board alone."""
    from alpha_zero_b200.ai.mcts import MCTS
    from alpha_zero_b200.ai.node import Node
    from alpha_zero_b200.games import TicTacToe
    board = np.zeros((3, 3, 3), np.float32)
    for (r, c), p in (((0, 0), 0), ((1, 1), 1), ((0, 1), 0), ((2, 2), 1)):      # X: (0,0) (0,1); O: (1,1) (2,2); X to move
        board[p, r, c] = 1
    before = board.copy()
    root = Node(None, None, 0, 4)
    np.random.seed(0)
    MCTS.mcts(None, board, root, TicTacToe, 600)
    assert np.array_equal(board, before)
    assert root.visit == 600 and sum(ch.visit for ch in root.children) == 599
    assert [ch.prevAction for ch in root.children] == TicTacToe.get_valid_moves(board)
    assert root.max_visit_child().prevAction == (0, 2)
    assert all(ch.prior == 0.0 for ch in root.children)
