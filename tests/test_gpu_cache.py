"""The device evaluation cache (MCTS.cache analogue, ai/mcts.py:7,37-51) must be TRANSPARENT: searches with the cache
on, off, warm or cold give bit-identical trees for a deterministic evaluator; it only removes evaluator rows.  Also the
life-cycle hazards around it: evaluator changes, forests destroyed while a net keeps captured step graphs."""
import numpy as np
import pytest

import az_oracle as orc
import c_oracle as co
from gpu_util import words_from_cells, random_positions, compare_forest_with_oracle

pytestmark = pytest.mark.gpu


def _children(f):
    return [(c["cells"].tolist(), c["visits"].tolist(), c["W"].tolist(), c["P"].tolist(), c["root_visit"], c["root_value"])
            for c in (f.root_children(t) for t in range(f.n_trees))]


@pytest.mark.parametrize("name,n,iters,vl", [("connect4", 96, 300, 1), ("gomoku", 64, 200, 1), ("gomoku", 64, 240, 8), ("tictactoe", 32, 100, 1)])
def test_cache_on_off_identical_and_equal_to_oracle(name, n, iters, vl):
    import alpha_zero_b200 as az
    g = {"tictactoe": orc.TICTACTOE, "connect4": orc.CONNECT4, "gomoku": orc.GOMOKU}[name]
    rng = np.random.default_rng(7)
    cells, players, mcs = random_positions(g, n, rng, 5)
    cells[n // 2:] = cells[:n - n // 2]                         # duplicate roots: leaves of different trees collide in one launch
    players[n // 2:] = players[:n - n // 2]
    mcs[n // 2:] = mcs[:n - n // 2]
    ref = co.search_batch(g, cells, players, mcs, iters, vl, seed=3, qbits=8)
    spec = az.game_spec(g.name, g.rows)
    on = az.Forest(spec, n, iters, vl=vl)
    off = az.Forest(spec, n, iters, vl=vl)
    off.cache_config(0)
    assert on.cache_info()["entries"] > 0 and off.cache_info()["entries"] == 0
    for f in (on, off):
        f.set_roots_words(words_from_cells(g, cells), players, mcs)
        f.stats(reset=True)
        f.search_hash(iters, 3, 8)
        compare_forest_with_oracle(f, ref)
    s_on, s_off = on.stats(), off.stats()
    assert s_off["cache_hits"] == 0 and s_on["cache_hits"] > 0
    for k in ("sims", "terminal", "expansions", "children", "selects", "child_reads"):
        assert s_on[k] == s_off[k], k
    # warm cache: the same search again is answered (almost) entirely from the cache and is still the same search
    on.stats(reset=True)
    on.search_hash(iters, 3, 8)
    compare_forest_with_oracle(on, ref)
    s_warm = on.stats()
    assert s_warm["cache_hits"] >= 0.9 * s_warm["expansions"]
    # another evaluator: cached answers of the old one must not be used
    ref2 = co.search_batch(g, cells, players, mcs, iters, vl, seed=4, qbits=8)
    on.search_hash(iters, 4, 8)
    compare_forest_with_oracle(on, ref2)
    # a tiny table (every window full most of the time) is still transparent
    on.cache_config(4)
    on.search_hash(iters, 3, 8)
    compare_forest_with_oracle(on, ref)
    on.close()
    off.close()


def test_successive_plies_reuse_the_previous_search():
    """The reference's 38-40 % hit rate comes from searching the position after the move just searched: most of the new
    root's subtree was evaluated one search earlier."""
    import alpha_zero_b200 as az
    g = orc.GOMOKU
    n, iters = 64, 400
    rng = np.random.default_rng(1)
    cells, players, mcs = random_positions(g, n, rng, 6)
    spec = az.game_spec(g.name, g.rows)
    f = az.Forest(spec, n, iters, vl=1)
    f.set_roots_words(words_from_cells(g, cells), players, mcs)
    f.search_hash(iters, 2, 12)
    vis, _ = f.root_visits()
    best = vis.argmax(axis=1)
    ok = np.ones(n, bool)
    for t in range(n):                                          # play the most visited move (skip games it ends)
        cells[t, best[t]] = players[t] + 1
        if co.winner(g, cells[t], players[t], best[t]) != -1:
            ok[t] = False
    players2, mcs2 = 1 - players, mcs + 1
    f.set_roots_words(words_from_cells(g, cells), players2, mcs2)
    f.set_active(ok.astype(np.uint8))
    f.stats(reset=True)
    f.search_hash(iters, 2, 12)
    st = f.stats()
    ref = co.search_batch(g, cells[ok], players2[ok], mcs2[ok], iters, 1, seed=2, qbits=12)
    idx = np.nonzero(ok)[0]
    for j, t in enumerate(idx[:16]):
        ch = f.root_children(int(t))
        nc = int(ref["n_children"][j])
        assert ch["visits"].tolist() == ref["visits"][j, :nc].tolist() and ch["W"].tolist() == ref["W"][j, :nc].tolist()
    assert st["cache_hits"] > 0.25 * st["expansions"], st
    f.close()


def test_net_search_with_cache_is_repeatable_and_replays_in_oracle():
    """Net-driven search: log every (position -> priors, value) the GPU produced step by step, replay in the C oracle;
    the CUDA-graph path (cache warm) must give the same tree."""
    import torch
    import alpha_zero_b200 as az
    from alpha_zero_b200.ai.nn import Net, B200Net
    g, n, iters = orc.CONNECT4, 96, 160
    rng = np.random.default_rng(5)
    torch.manual_seed(3)
    net = Net((g.rows, g.cols), 5, 128, g.action_dim, 8, 1, g.planes, dropout=0.0).eval()
    eng = B200Net(net, max_batch=n)
    cells, players, mcs = random_positions(g, n, rng, 6)
    spec = az.game_spec(g.name, g.rows)
    f = az.Forest(spec, n, iters, vl=1)
    f.set_roots_words(words_from_cells(g, cells), players, mcs)
    f.begin_move(iters)
    keys, pri, val = [], [], []
    for _ in range(2 * iters + 8):
        f.advance()
        eng.eval_leaves(f)
        r = f.read_leaves()
        if r["n"]:
            k = np.zeros((r["n"], 3), np.uint64)
            k[:, :2] = r["words"].reshape(r["n"], -1)
            k[:, 2] = r["players"]
            keys.append(k)
            pri.append(r["priors"])
            val.append(r["values"])
        if f.counts()[1] >= n:
            break
    assert f.counts()[1] == n
    st = f.stats()
    assert sum(len(k) for k in keys) == st["expansions"] - st["cache_hits"]      # one evaluator row per miss, none per hit
    table = co.Table(np.concatenate(keys), np.concatenate(pri), np.concatenate(val))
    ref = co.search_batch(g, cells, players, mcs, iters, 1, table=table)
    assert int(ref["stats"][7]) == 0
    compare_forest_with_oracle(f, ref)
    f.search_net(eng, iters)                                    # graph path, warm cache
    compare_forest_with_oracle(f, ref)
    f.close()
    eng.close()


def test_forest_destroyed_and_recreated_under_a_live_net():
    """the step graph bakes the forest's device pointers in; a new forest at the freed address with the same
    iters must not replay the stale graph."""
    import torch
    import alpha_zero_b200 as az
    from alpha_zero_b200.ai.nn import Net, B200Net
    g, iters = orc.CONNECT4, 64
    torch.manual_seed(4)
    net = Net((g.rows, g.cols), 5, 128, g.action_dim, 8, 1, g.planes, dropout=0.0).eval()
    eng = B200Net(net, max_batch=256)
    spec = az.game_spec(g.name, g.rows)
    rng = np.random.default_rng(9)
    results = []
    for n in (128, 40, 128, 200):
        cells, players, mcs = random_positions(g, n, rng, 6)
        f = az.Forest(spec, n, iters, vl=1)
        f.set_roots_words(words_from_cells(g, cells), players, mcs)
        f.search_net(eng, iters)
        vis, _ = f.root_visits()
        assert (vis.sum(axis=1) == iters - 1).all()
        f2 = az.Forest(spec, n, iters, vl=1)                    # a second live forest on the same net keeps its own graph
        f2.cache_config(0)
        f2.set_roots_words(words_from_cells(g, cells), players, mcs)
        f2.search_net(eng, iters)
        vis2, _ = f2.root_visits()
        assert (vis2.sum(axis=1) == iters - 1).all()
        f.search_net(eng, iters)                                # and the first one still works afterwards
        vis3, _ = f.root_visits()
        assert np.array_equal(vis, vis3)
        results.append((vis, vis2))
        f.close()
        f2.close()
    eng.close()
