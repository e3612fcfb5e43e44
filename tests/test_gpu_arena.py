"""Batched arena (alpha_zero_b200/arena.py) against the oracle: every search of every game is replayed in the C oracle
with the evaluator / simulation budget of the side to move and the logged Dirichlet noise; the moves must be the first
most-visited child in child order (or np.random.choice's pick for the logged uniform), the winners must follow from the
oracle's rules, and compare() must return what the reference's sequential loop (test.py:106-140) returns for those
outcomes, early stopping included."""
import numpy as np
import pytest

import az_oracle as orc
import c_oracle as co

pytestmark = pytest.mark.gpu


def game_cls(g):
    from alpha_zero_b200 import games
    G = {"tictactoe": games.TicTacToe, "connect4": games.Connect4, "gomoku": games.Gomoku}[g.name]
    if g.name == "gomoku":
        G.rows = G.cols = g.rows
        G.action_dim = G.state_dim = g.rows * g.cols
    return G


def reference_compare_result(winners, iterations, early_stopping):
    """test.py:106-140 given each game's winner (0 / 1 / -1)."""
    wc = [0, 0, 0]
    for i in range(iterations):
        first = i < iterations // 2
        if winners[i] == 0:
            wc[0 if first else 1] += 1
        elif winners[i] == 1:
            wc[1 if first else 0] += 1
        else:
            wc[0] += 0.5
            wc[1] += 0.5
            wc[2] += 1
        if early_stopping:
            if wc[1] >= int(iterations * 0.55):
                return 1
            if wc[1] + iterations - (i + 1) < int(iterations * 0.55):
                return 0
    return wc[1] / iterations


@pytest.mark.parametrize("name,iterations,sampling,qbits", [("gomoku", 6, False, 10), ("connect4", 7, True, 12), ("gomoku", 4, False, 2)])
def test_compare_replays_in_oracle(name, iterations, sampling, qbits):
    from alpha_zero_b200 import arena
    from alpha_zero_b200.ai import HashEvaluator
    g = {"connect4": orc.CONNECT4, "gomoku": orc.GOMOKU}[name]
    G = game_cls(g)
    evs = (HashEvaluator(21, qbits), HashEvaluator(22, qbits))
    iters = (40, 28)
    log = []
    got = arena.compare(G, evs[0], evs[1], iters[0], iters[1], iterations, sampling, early_stopping=False,
                        rng=np.random.RandomState(3), log=log)
    ncell = g.rows * g.cols
    cells = np.zeros((iterations, ncell), np.uint8)
    winners = np.full(iterations, -2)
    first_is_best = np.arange(iterations) < iterations // 2
    for ent in log:
        ply = ent["ply"]
        for j, i in enumerate(ent["live"]):
            which = 0 if first_is_best[i] == (ply % 2 == 0) else 1
            ref = co.search_batch(g, cells[i:i + 1], np.array([ply & 1], np.int32), np.array([ply], np.int32), iters[which], 1,
                                  seed=evs[which].seed, qbits=qbits, noise=ent["noise"][j:j + 1])
            nc = int(ref["n_children"][0])
            ch_cells, ch_vis = ref["cells"][0, :nc], ref["visits"][0, :nc]
            ai = ch_cells % g.cols if g.gid == 1 else ch_cells
            vis = np.zeros(g.action_dim, np.int64)
            vis[ai] = ch_vis
            assert np.array_equal(ent["visits"][j], vis), f"game {i} ply {ply}: visits"
            cell = int(ch_cells[int(np.argmax(ch_vis))])             # first maximum in child order
            if ent["u"][j] >= 0:
                p = vis / np.sum(vis.astype(np.float64))
                cdf = p.cumsum()
                cdf /= cdf[-1]
                a = int(cdf.searchsorted(ent["u"][j], side="right"))
                cell = int(ch_cells[list(ai).index(a)])
            assert int(ent["cells"][j]) == cell, f"game {i} ply {ply}: move"
            cells[i, cell] = (ply & 1) + 1
            w = co.winner(g, cells[i], ply & 1, cell)
            if w != -1 or ply + 1 == ncell:
                winners[i] = w
    assert (winners != -2).all()
    assert got == reference_compare_result(winners, iterations, False)
    assert sampling == any((e["u"] >= 0).any() for e in log)


def test_early_stopping_rule_and_compete():
    from alpha_zero_b200 import arena
    from alpha_zero_b200.ai import HashEvaluator
    g = orc.GOMOKU
    G = game_cls(g)
    a, b = HashEvaluator(5, 10), HashEvaluator(6, 10)
    np.random.seed(1)
    r = arena.compare(G, a, b, 32, 32, 8, False, early_stopping=True)
    assert r in (0, 1) or 0.0 <= r <= 1.0
    w, board = arena.compete(G, a, b, 24, 24)
    assert w in (-1, 0, 1) and board.shape == (2, g.rows, g.cols) and board.sum() >= 9
    for winners in ([1, 1, 1, 1, 0, 0, 0, 0], [0, 0, 0, 0, 1, 1, 1, 1], [-1] * 8, [0, 1, 0, 1, 0, 1, 0, 1]):
        assert reference_compare_result(winners, 8, True) in (0, 1)


def test_vanilla_opponent_runs():
    """main.py:74-76 plays the net against vanilla MCTS (model=None: UCB1 + rollouts, no Dirichlet noise)."""
    from alpha_zero_b200 import arena
    from alpha_zero_b200.ai import HashEvaluator
    G = game_cls(orc.CONNECT4)
    np.random.seed(2)
    rate = arena.compare(G, None, HashEvaluator(3, 12), 40, 40, 4, False, early_stopping=False)
    assert 0.0 <= rate <= 1.0
