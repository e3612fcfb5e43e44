"""ctypes binding of oracle/libaz_oracle.so (the C oracle).  TEST INFRASTRUCTURE: only
tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs import this."""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build():
    subprocess.check_call(["make", "-s", "-C", HERE])


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(HERE, "libaz_oracle.so")
        if not os.path.exists(path) or os.path.getmtime(path) < os.path.getmtime(os.path.join(HERE, "az_oracle.c")):
            build()
        L = C.CDLL(path)
        L.azo_table_create.restype = C.c_void_p
        L.azo_table_create.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        L.azo_table_destroy.argtypes = [C.c_void_p]
        L.azo_search_batch.restype = C.c_int
        L.azo_search_batch.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                       C.c_int, C.c_int, C.c_uint64, C.c_int, C.c_void_p, C.c_void_p, C.c_int] + \
                                      [C.c_void_p] * 8
        L.azo_legal.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        L.azo_winner.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int]
        L.azo_puct.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        _LIB = L
    return _LIB


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def cells_from_planes(board):
    """f32 [C,H,W] reference board -> uint8 cells (0 empty, 1 side 0, 2 side 1)."""
    b = np.asarray(board)
    return (b[0] == 1).astype(np.uint8).reshape(-1) + 2 * (b[1] == 1).astype(np.uint8).reshape(-1)


class Table:
    """(canonical words, side-to-move) -> (priors, value) evaluator table (MCTS.cache preload route)."""

    def __init__(self, keys, priors, values):
        self.keys = np.ascontiguousarray(keys, dtype=np.uint64)
        self.priors = np.ascontiguousarray(priors, dtype=np.float32)
        self.values = np.ascontiguousarray(values, dtype=np.float32)
        n, kl = self.keys.shape
        self.h = lib().azo_table_create(n, (kl - 1) // 2, self.priors.shape[1], _p(self.keys), _p(self.priors), _p(self.values))

    def __del__(self):
        if getattr(self, "h", None):
            lib().azo_table_destroy(self.h)
            self.h = None


def search_batch(g, cells, players, mcs, iters, vl=1, seed=0, qbits=16, noise=None, table=None):
    """n independent searches.  cells uint8 [n][rows*cols].  Returns dict of arrays."""
    cells = np.ascontiguousarray(cells, dtype=np.uint8)
    n = cells.shape[0]
    players = np.ascontiguousarray(players, dtype=np.int32)
    mcs = np.ascontiguousarray(mcs, dtype=np.int32)
    maxch = 9 if g.gid == 0 else 7 if g.gid == 1 else g.rows * g.cols
    nc = np.zeros(n, np.int32)
    oc = np.full((n, maxch), -1, np.int32)
    ov = np.zeros((n, maxch), np.int32)
    oW = np.zeros((n, maxch), np.float64)
    oP = np.zeros((n, maxch), np.float64)
    rv = np.zeros(n, np.int32)
    rW = np.zeros(n, np.float64)
    stats = np.zeros(8, np.uint64)
    if noise is not None:
        noise = np.ascontiguousarray(noise, dtype=np.float64)
    rc = lib().azo_search_batch(g.gid, g.rows, g.cols, n, _p(cells), _p(players), _p(mcs), iters, vl, seed, qbits,
                                _p(noise), table.h if table is not None else None, maxch,
                                _p(nc), _p(oc), _p(ov), _p(oW), _p(oP), _p(rv), _p(rW), _p(stats))
    if rc != 0:
        raise RuntimeError(f"azo_search_batch failed: {rc}")
    return dict(n_children=nc, cells=oc, visits=ov, W=oW, P=oP, root_visit=rv, root_value=rW, stats=stats)


def set_ucb1(on):
    """Vanilla-MCTS selection (utils.py:36,43) instead of PUCT for the following searches."""
    lib().azo_set_ucb1(int(bool(on)))


def legal(g, cells):
    out = np.zeros(g.rows * g.cols, np.int32)
    c = np.ascontiguousarray(cells, dtype=np.uint8)
    n = lib().azo_legal(g.gid, g.rows, g.cols, _p(c), _p(out))
    return out[:n].tolist()


def winner(g, cells, player, cell):
    c = np.ascontiguousarray(cells, dtype=np.uint8)
    return lib().azo_winner(g.gid, g.rows, g.cols, _p(c), int(player), int(cell))


def puct(parent_visit, is_f64, priors, visits, W):
    priors = np.ascontiguousarray(priors, np.float64)
    visits = np.ascontiguousarray(visits, np.int32)
    W = np.ascontiguousarray(W, np.float64)
    out = np.zeros(len(priors), np.float64)
    sel = lib().azo_puct(len(priors), parent_visit, int(is_f64), _p(priors), _p(visits), _p(W), _p(out))
    return sel, out
