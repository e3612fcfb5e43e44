"""Shared helpers of the GPU parity tests."""
import numpy as np

import az_oracle as orc
import c_oracle as co


def has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def words_from_cells(g, cells):
    """uint8 [n][cells] (0/1/2) -> uint64 [n][2][nw]."""
    cells = np.asarray(cells, dtype=np.uint8)
    n = cells.shape[0]
    nw = 1 if g.rows * g.cols <= 64 else 4               # the C ABI's word count per side
    out = np.zeros((n, 2, nw), dtype=np.uint64)
    for side in (0, 1):
        m = (cells == side + 1)
        for w in range(nw):
            seg = m[:, w * 64:(w + 1) * 64].astype(np.uint64)
            out[:, side, w] = (seg << np.arange(seg.shape[1], dtype=np.uint64)).sum(axis=1, dtype=np.uint64)
    return out


def random_positions(g, n, rng, max_plies):
    """n unfinished positions after 0..max_plies random legal plies (reference order rules
    through the C oracle).  Returns cells uint8 [n][cells], players, move_counts."""
    ncell = g.rows * g.cols
    cells = np.zeros((n, ncell), np.uint8)
    players = np.zeros(n, np.int32)
    mcs = np.zeros(n, np.int32)
    for t in range(n):
        target = int(rng.integers(0, max_plies + 1)) if max_plies else 0
        while True:
            b = np.zeros(ncell, np.uint8)
            pl, ok = 0, True
            for _ in range(target):
                lm = co.legal(g, b)
                c = lm[int(rng.integers(len(lm)))]
                b[c] = pl + 1
                if co.winner(g, b, pl, c) != -1:
                    ok = False
                    break
                pl = 1 - pl
            if ok:
                break
        cells[t], players[t], mcs[t] = b, pl, target
    return cells, players, mcs


def compare_forest_with_oracle(forest, ref, trees=None):
    """Bit-exact comparison of root children of the GPU forest with the C oracle result."""
    n = forest.n_trees
    idx = range(n) if trees is None else trees
    vis, q = forest.root_visits()
    g = forest.g
    for t in idx:
        nc = int(ref["n_children"][t])
        ch = forest.root_children(t)
        assert ch["cells"].tolist() == ref["cells"][t, :nc].tolist(), f"tree {t}: child order"
        assert ch["visits"].tolist() == ref["visits"][t, :nc].tolist(), f"tree {t}: visits"
        assert ch["W"].tolist() == ref["W"][t, :nc].tolist(), f"tree {t}: W"
        assert ch["P"].tolist() == ref["P"][t, :nc].tolist(), f"tree {t}: priors"
        assert ch["root_visit"] == int(ref["root_visit"][t])
        assert ch["root_value"] == float(ref["root_value"][t])
    # visit-count policy in action order for ALL trees
    exp = np.zeros_like(vis)
    for t in range(n):
        nc = int(ref["n_children"][t])
        cells = ref["cells"][t, :nc]
        ai = cells % g.cols if g.gid == 1 else cells
        exp[t, ai] = ref["visits"][t, :nc]
    assert np.array_equal(vis, exp)
    rq = np.where(ref["root_visit"] > 0, ref["root_value"] / np.maximum(ref["root_visit"], 1), 0.0)
    assert np.array_equal(q, rq)
