"""Exact-replay parity at the BENCHMARKED configurations.

The searches are driven by the tensor-core net at the shapes bench.py measures; every (position -> priors, value) answer
the GPU produced is logged step by step, the C oracle replays sampled trees from that table (the MCTS.cache preload
route of SURVEY section 8c) and must reproduce child order, visit counts, W (f64) and priors bit for bit.  The CUDA-graph
path (az_forest_search_net: graph replay + PDL + evaluation cache) must then give the same trees as the eager steps, on
a warm and on a cold cache.  Last: what bf16 evaluation does to the SEARCH -- visit-count policies of a B200-net search
against the same search driven by the torch fp32 Net."""
import numpy as np
import pytest

import az_oracle as orc
import c_oracle as co
from gpu_util import words_from_cells, random_positions

pytestmark = pytest.mark.gpu


def _logged_search(f, eng, iters, n):
    """Eager steps with every evaluation batch read back.  Returns the evaluator table of the whole search."""
    nw = f.g.nw
    f.begin_move(iters)
    keys, pri, val = [], [], []
    for _ in range(2 * iters + 8):
        f.advance()
        eng.eval_leaves(f)
        r = f.read_leaves()
        if r["n"]:
            k = np.zeros((r["n"], 2 * nw + 1), np.uint64)
            k[:, :2 * nw] = r["words"].reshape(r["n"], -1)
            if f.g.planes == 3:                                   # 2-plane games: the canonical words are the whole key
                k[:, -1] = r["players"]
            keys.append(k)
            pri.append(r["priors"].copy())
            val.append(r["values"].copy())
        if f.counts()[1] >= n:
            break
    assert f.counts()[1] == n, "search did not finish"
    return co.Table(np.concatenate(keys), np.concatenate(pri), np.concatenate(val)), sum(len(k) for k in keys)


def _check_sample(f, g, ref, sample):
    for j, t in enumerate(sample):
        nc = int(ref["n_children"][j])
        ch = f.root_children(int(t))
        assert ch["cells"].tolist() == ref["cells"][j, :nc].tolist(), f"tree {t}: child order"
        assert ch["visits"].tolist() == ref["visits"][j, :nc].tolist(), f"tree {t}: visits"
        assert ch["W"].tolist() == ref["W"][j, :nc].tolist(), f"tree {t}: W"
        assert ch["P"].tolist() == ref["P"][j, :nc].tolist(), f"tree {t}: priors"
        assert ch["root_visit"] == int(ref["root_visit"][j]) and ch["root_value"] == float(ref["root_value"][j])


def _replay(name, rows, n, iters, vl, dim, depth, patch, n_sample, max_plies, want_flags=0, graph=True):
    import torch
    import alpha_zero_b200 as az
    from alpha_zero_b200.ai.nn import Net, B200Net
    g = {"connect4": orc.CONNECT4}.get(name) or orc.gomoku_spec(rows)
    rng = np.random.default_rng(17)
    torch.manual_seed(0)
    net = Net((g.rows, g.cols), patch, dim, g.action_dim, 8, depth, g.planes, dropout=0.1).eval()
    eng = B200Net(net, max_batch=n * vl)
    assert eng.path_flags() & want_flags == want_flags, f"evaluator path flags {eng.path_flags()}"
    base, bp, bm = random_positions(g, 256, rng, max_plies)           # 256 distinct openings, tiled over the trees
    idx = np.arange(n) % 256
    cells, players, mcs = base[idx], bp[idx], bm[idx]
    spec = az.game_spec(g.name, g.rows)
    f = az.Forest(spec, n, iters, vl=vl)
    f.set_roots_words(words_from_cells(g, cells), players, mcs)
    f.stats(reset=True)
    table, rows_eval = _logged_search(f, eng, iters, n)
    st = f.stats()
    assert st["sims"] == n * iters
    assert rows_eval == st["expansions"] - st["cache_hits"]           # one evaluator row per miss, none per hit
    sample = rng.choice(n, size=n_sample, replace=False)
    ref = co.search_batch(g, cells[sample], players[sample], mcs[sample], iters, vl, table=table)
    assert int(ref["stats"][7]) == 0, "the oracle reached a position the GPU never evaluated"
    _check_sample(f, g, ref, sample)
    vis_eager, q_eager = f.root_visits()
    if graph:
        for cold in (False, True):                                    # graph replay + PDL: warm cache, then cold
            if cold:
                f.cache_clear()
            f.search_net(eng, iters)
            vis, q = f.root_visits()
            assert np.array_equal(vis, vis_eager) and np.array_equal(q, q_eager), f"graph path differs (cold={cold})"
        _check_sample(f, g, ref, sample[:16])
    f.close()
    eng.close()
    return st


def test_connect4_benchmark_config():
    """BASELINE.json configs[1]: Connect4 6x7, 4096 trees x 800 sims, vl 1, Net D=512 patch 5 depth 1 (main.py:134)."""
    st = _replay("connect4", 6, 4096, 800, 1, 512, 1, 5, 64, 7, want_flags=2 | 8 | 16)
    assert st["cache_hits"] > 0


def test_gomoku_benchmark_config_cta_pair_gemms():
    """configs[2] shape: Gomoku 7x7, 8192 trees, vl 8, engine sized for 65536 rows -> CTA-pair cls-row GEMMs at their default
    threshold (fewer simulations per move than the bench: the table of a full 800-sim search would not fit the test)."""
    _replay("gomoku", 7, 8192, 160, 8, 512, 1, 5, 64, 10, want_flags=2 | 4 | 8 | 16)


def test_gomoku_15x15_generic_path():
    _replay("gomoku", 15, 96, 120, 1, 128, 1, 5, 96, 8, graph=True)


def test_depth2_net_of_main_py_187():
    """main.py:187-188: D=256, depth 2, patch 5 on 7x7 (generic path: full blocks before the cls-row tail)."""
    _replay("gomoku", 7, 256, 160, 1, 256, 2, 5, 64, 10, graph=True)


def test_bf16_search_against_fp32_search():
    """north_star asks Q / pi within 1e-5 with a FIXED evaluator; the bf16 net is a different evaluator than torch fp32, so
    the honest number is what that does to the search.  Connect4, 256 roots x 200 sims: visit-count policies of the B200-net
    search vs the same search driven by the torch fp32 Net (external-evaluator path).  Figures are printed for DESIGN.md."""
    import torch
    import alpha_zero_b200 as az
    from alpha_zero_b200.ai import MCTS
    from alpha_zero_b200.ai.nn import Net, B200Net
    g = orc.CONNECT4
    n, iters = 256, 200
    rng = np.random.default_rng(23)
    torch.manual_seed(0)
    net = Net((g.rows, g.cols), 5, 512, g.action_dim, 8, 1, g.planes, dropout=0.1).eval()
    eng = B200Net(net, max_batch=n)
    cells, players, mcs = random_positions(g, n, rng, 10)
    spec = az.game_spec(g.name, g.rows)
    f = az.Forest(spec, n, iters, vl=1)
    f.set_roots_words(words_from_cells(g, cells), players, mcs)
    f.search_net(eng, iters)
    vis16, q16 = f.root_visits()
    gpu_net = Net((g.rows, g.cols), 5, 512, g.action_dim, 8, 1, g.planes, dropout=0.1)
    gpu_net.load_state_dict(net.state_dict())
    gpu_net = gpu_net.cuda().eval()
    f.set_roots_words(words_from_cells(g, cells), players, mcs)
    MCTS._search_external(f, gpu_net, iters)
    vis32, q32 = f.root_visits()
    f.close()
    eng.close()
    pi16, pi32 = vis16 / vis16.sum(1, keepdims=True), vis32 / vis32.sum(1, keepdims=True)
    top1 = float((pi16.argmax(1) == pi32.argmax(1)).mean())
    dpi = np.abs(pi16 - pi32)
    kl = float(np.mean(np.sum(np.where(pi32 > 0, pi32 * np.log(np.maximum(pi32, 1e-12) / np.maximum(pi16, 1e-12)), 0.0), axis=1)))
    print(f"\nbf16 vs fp32 search, Connect4 {n} roots x {iters} sims: top-1 agreement {top1:.4f}, max|dpi| {dpi.max():.4f}, "
          f"mean|dpi| {dpi.mean():.5f}, mean KL(fp32||bf16) {kl:.5f}, max|dQ| {np.abs(q16 - q32).max():.4f}, identical roots "
          f"{float((vis16 == vis32).all(1).mean()):.3f}")
    assert top1 >= 0.95 and dpi.mean() <= 0.005            # measured on B200: 0.9961 and 0.00068
