"""GPU parity of the forest kernels (through the C ABI) against the reference's golden
outputs and against the C oracle.  Integer/index work and f64 sums: bit-exact."""
import numpy as np
import pytest

import az_oracle as orc
import c_oracle as co
from golden_util import load_json, spec_of, board_from_words, position_after, fhex
from gpu_util import words_from_cells, random_positions, compare_forest_with_oracle

pytestmark = pytest.mark.gpu


def az():
    import alpha_zero_b200
    return alpha_zero_b200


def spec(g):
    return az().game_spec(g.name, g.rows)


@pytest.mark.parametrize("name,g", [("rules_tictactoe.json", orc.TICTACTOE), ("rules_connect4.json", orc.CONNECT4),
                                    ("rules_gomoku7.json", orc.GOMOKU), ("rules_gomoku15.json", orc.gomoku_spec(15))])
def test_rules_golden(name, g):
    from alpha_zero_b200.forest import rules_legal, rules_winner
    cells, legal, after, players, moves, winners = [], [], [], [], [], []
    for plies in load_json(name):
        for rec in plies:
            b = board_from_words(g, rec["p0"], rec["p1"])
            c = co.cells_from_planes(b)
            cells.append(c.copy())
            legal.append(rec["legal"])
            mv = rec["move"][0] * g.cols + rec["move"][1]
            c[mv] = rec["player"] + 1
            after.append(c)
            players.append(rec["player"])
            moves.append(mv)
            winners.append(rec["winner"])
    got = rules_legal(spec(g), words_from_cells(g, np.stack(cells)))
    assert got == legal
    w = rules_winner(spec(g), words_from_cells(g, np.stack(after)), players, moves)
    assert w.tolist() == winners


@pytest.mark.parametrize("g,n", [(orc.TICTACTOE, 20000), (orc.CONNECT4, 100000), (orc.GOMOKU, 100000),
                                 (orc.gomoku_spec(15), 20000), (orc.gomoku_spec(9), 20000)])
def test_rules_random_differential(g, n):
    """Random (not necessarily reachable) boards: legal list incl. order, and winner through
    a random stone, GPU vs C oracle."""
    from alpha_zero_b200.forest import rules_legal, rules_winner
    rng = np.random.default_rng(11)
    ncell = g.rows * g.cols
    dens = rng.random((n, 1))
    r = rng.random((n, ncell))
    cells = np.where(r < dens * 0.5, 1, np.where(r < dens, 2, 0)).astype(np.uint8)
    if g.gid == 1:                                           # gravity: stack stones at the bottom
        cc = cells.reshape(n, g.rows, g.cols)
        for t in range(n):
            for c in range(g.cols):
                col = cc[t, :, c]
                st = col[col > 0]
                col[:] = 0
                if len(st):
                    col[g.rows - len(st):] = st
        cells = cc.reshape(n, ncell)
    full = (cells > 0).all(axis=1)
    cells[full, 0] = 0
    words = words_from_cells(g, cells)
    got = rules_legal(spec(g), words)
    sample = rng.choice(n, size=min(n, 4000), replace=False)
    for t in sample:
        assert got[t] == co.legal(g, cells[t]), t
    players = rng.integers(0, 2, n).astype(np.int32)
    mv = np.zeros(n, np.int32)
    for t in range(n):
        own = np.nonzero(cells[t] == players[t] + 1)[0]
        if len(own) == 0:
            cells[t, 0] = players[t] + 1
            own = np.array([0])
        mv[t] = own[int(rng.integers(len(own)))]
    w = rules_winner(spec(g), words_from_cells(g, cells), players, mv)
    for t in sample:
        assert w[t] == co.winner(g, cells[t], int(players[t]), int(mv[t])), t
    assert (w != -1).sum() > 0


def test_search_golden_cases():
    """Every search case recorded from the unmodified reference (tests/golden/search.json):
    child order, visit counts, W (f64) and priors bit-exact, incl. Dirichlet-noise roots."""
    forests = {}
    cases = load_json("search.json")
    for case in cases:
        g = spec_of(case)
        key = (g.name, g.rows)
        if key not in forests:
            forests[key] = az().Forest(spec(g), 1, 800)
        f = forests[key]
        board, player, mc = position_after(g, case["opening"])
        f.set_roots_planes(board[None], [player], [mc])
        f.set_noise(None if case["noise"] is None else np.asarray([case["noise"]]))
        f.search_hash(case["iters"], case["seed"], case["qbits"])
        ch, ref = f.root_children(0), case["result"]
        assert ch["cells"].tolist() == [m[0] * g.cols + m[1] for m in ref["moves"]]
        assert ch["visits"].tolist() == ref["visits"]
        assert ch["W"].tolist() == [fhex(v) for v in ref["values"]]
        assert ch["P"].tolist() == [fhex(v) for v in ref["priors"]]
        assert ch["root_visit"] == ref["root_visit"] == case["iters"]
        assert ch["root_value"] == fhex(ref["root_value"])
    for f in forests.values():
        f.close()


@pytest.mark.parametrize("g,n,iters,vl,plies", [
    (orc.CONNECT4, 4096, 800, 1, 7),          # BASELINE config 2: bit-exactness at full size
    (orc.GOMOKU, 2048, 800, 1, 12),
    (orc.TICTACTOE, 512, 100, 1, 5),          # BASELINE config 1 shape
    (orc.gomoku_spec(15), 128, 200, 1, 30),
    (orc.GOMOKU, 1024, 800, 8, 12),           # BASELINE config 3: virtual-loss batch 8
    (orc.CONNECT4, 1024, 400, 4, 7),
])
def test_forest_vs_c_oracle(g, n, iters, vl, plies):
    rng = np.random.default_rng(1234)
    cells, players, mcs = random_positions(g, n, rng, plies)
    ref = co.search_batch(g, cells, players, mcs, iters, vl, seed=0, qbits=16)
    f = az().Forest(spec(g), n, iters, vl=vl)
    f.set_roots_words(words_from_cells(g, cells), players, mcs)
    f.search_hash(iters, 0, 16)
    compare_forest_with_oracle(f, ref, trees=range(0, n, max(1, n // 64)))
    st = f.stats()
    assert st["sims"] == n * iters
    assert st["sims"] == int(ref["stats"][0]) and st["expansions"] == int(ref["stats"][2])
    assert st["children"] == int(ref["stats"][3]) and st["child_reads"] == int(ref["stats"][5])
    f.close()


def test_noise_and_inactive_trees():
    g = orc.GOMOKU
    rng = np.random.default_rng(5)
    n, iters = 256, 200
    cells, players, mcs = random_positions(g, n, rng, 10)
    noise = rng.dirichlet([0.03] * g.action_dim, size=n)
    ref = co.search_batch(g, cells, players, mcs, iters, 1, seed=9, qbits=4, noise=noise)
    f = az().Forest(spec(g), n, iters)
    f.set_roots_words(words_from_cells(g, cells), players, mcs)
    f.set_noise(noise)
    active = np.ones(n, np.uint8)
    active[::7] = 0
    f.set_active(active)
    f.search_hash(iters, 9, 4)
    for t in range(0, n, 5):
        ch = f.root_children(t)
        if not active[t]:
            assert ch["root_visit"] == 0
            continue
        nc = int(ref["n_children"][t])
        assert ch["visits"].tolist() == ref["visits"][t, :nc].tolist()
        assert ch["W"].tolist() == ref["W"][t, :nc].tolist()
        assert ch["P"].tolist() == ref["P"][t, :nc].tolist()
    f.close()
