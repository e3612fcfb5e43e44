"""Forest: thousands of concurrent search trees in HBM, driven through the C ABI.

Host-side mirror of what the reference does with Python objects:
  Node tree            ai/node.py:20-40      -> flat arrays in HBM (csrc/az_forest.h)
  MCTS.mcts            ai/mcts.py:11-60      -> Forest.search_*()
  visit-count policy   utils.py:46-55        -> Forest.root_visits()
"""
import ctypes as C
from dataclasses import dataclass

import numpy as np

from . import lib as _lib


@dataclass(frozen=True)
class GameSpec:
    name: str
    gid: int
    rows: int
    cols: int
    planes: int
    k: int
    action_dim: int

    @property
    def state_dim(self):
        return self.rows * self.cols

    @property
    def nw(self):
        return 1 if self.rows * self.cols <= 64 else 4      # csrc instantiates 1 and 4 words

    @property
    def max_children(self):
        return {0: 9, 1: 7}.get(self.gid, self.rows * self.cols)


GAMES = {
    "tictactoe": GameSpec("tictactoe", 0, 3, 3, 3, 3, 9),      # games/tictactoe.py:10-12
    "connect4": GameSpec("connect4", 1, 6, 7, 3, 4, 7),        # games/connect4.py:8-10
    "gomoku": GameSpec("gomoku", 2, 7, 7, 2, 5, 49),           # games/gomoku.py:10-13
}


def game_spec(name, rows=None):
    g = GAMES[name]
    if name == "gomoku" and rows is not None and rows != g.rows:
        return GameSpec("gomoku", 2, rows, rows, 2, 5, rows * rows)
    return g


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def planes_to_words(g, planes):
    """f32 [n][C][H][W] -> uint64 [n][2][nw] side bitboards (bit r*cols+c)."""
    planes = np.asarray(planes)
    n = planes.shape[0]
    out = np.zeros((n, 2, g.nw), dtype=np.uint64)
    flat = (planes[:, :2].reshape(n, 2, -1) == 1)
    for w in range(g.nw):
        seg = flat[:, :, w * 64:(w + 1) * 64]
        weights = (np.uint64(1) << np.arange(seg.shape[2], dtype=np.uint64))
        out[:, :, w] = (seg.astype(np.uint64) * weights).sum(axis=2, dtype=np.uint64)
    return out


class Forest:
    """n_trees independent searches on one GPU.  `stream` is a raw cudaStream_t (int) or 0."""

    def __init__(self, game, n_trees, max_iters, vl=1, rows=None, stream=0):
        self.g = game if isinstance(game, GameSpec) else game_spec(game, rows)
        self.n_trees, self.max_iters, self.vl = int(n_trees), int(max_iters), int(vl)
        self.stream = stream
        self._L = _lib.load()
        h = C.c_void_p()
        _lib.check(self._L.az_forest_create(self.g.gid, self.g.rows, self.g.cols, self.n_trees, self.max_iters,
                                            self.vl, C.byref(h)), "az_forest_create")
        self.h = h

    def close(self):
        if getattr(self, "h", None):
            self._L.az_forest_destroy(self.h)
            self.h = None

    __del__ = close

    # ---- roots ------------------------------------------------------------------------
    def set_roots_planes(self, planes, players, move_counts):
        p = np.ascontiguousarray(planes, dtype=np.float32)
        assert p.shape == (self.n_trees, self.g.planes, self.g.rows, self.g.cols), p.shape
        pl = np.ascontiguousarray(players, dtype=np.int32)
        mc = np.ascontiguousarray(move_counts, dtype=np.int32)
        _lib.check(self._L.az_forest_set_roots_planes(self.h, _ptr(p), _ptr(pl), _ptr(mc), self.stream), "set_roots_planes")
        self._keep = (p, pl, mc)

    def set_roots_words(self, words, players, move_counts):
        w = np.ascontiguousarray(words, dtype=np.uint64).reshape(self.n_trees, 2, self.g.nw)
        pl = np.ascontiguousarray(players, dtype=np.int32)
        mc = np.ascontiguousarray(move_counts, dtype=np.int32)
        _lib.check(self._L.az_forest_set_roots_words(self.h, _ptr(w), _ptr(pl), _ptr(mc), self.stream), "set_roots_words")
        self._keep = (w, pl, mc)

    def set_active(self, active=None):
        a = None if active is None else np.ascontiguousarray(active, dtype=np.uint8)
        _lib.check(self._L.az_forest_set_active(self.h, _ptr(a), self.stream), "set_active")
        self._keep_a = a

    def set_noise(self, noise=None):
        """noise: f64 [n_trees][A] (the np.random.dirichlet draw of utils.py:21) or None."""
        nz = None if noise is None else np.ascontiguousarray(noise, dtype=np.float64)
        if nz is not None:
            assert nz.shape == (self.n_trees, self.g.action_dim)
        _lib.check(self._L.az_forest_set_noise(self.h, _ptr(nz), self.stream), "set_noise")
        self._keep_n = nz

    # ---- search -------------------------------------------------------------------------
    def begin_move(self, iters):
        _lib.check(self._L.az_forest_begin_move(self.h, int(iters), self.stream), "begin_move")

    def advance(self):
        _lib.check(self._L.az_forest_advance(self.h, self.stream), "advance")

    def counts(self):
        a, b = C.c_int(), C.c_int()
        _lib.check(self._L.az_forest_counts(self.h, C.byref(a), C.byref(b), self.stream), "counts")
        return a.value, b.value

    def eval_hash(self, seed, qbits=16):
        _lib.check(self._L.az_eval_hash(self.h, C.c_uint64(seed), int(qbits), self.stream), "eval_hash")

    def search_hash(self, iters, seed, qbits=16):
        _lib.check(self._L.az_forest_search_hash(self.h, int(iters), C.c_uint64(seed), int(qbits), self.stream),
                   "search_hash")

    def set_mode(self, ucb1):
        """True: vanilla MCTS (model=None) -- UCB1 selection, children without priors; False: PUCT."""
        _lib.check(self._L.az_forest_set_mode(self.h, 1 if ucb1 else 0), "set_mode")

    def eval_rollout(self, seed, counter=0):
        _lib.check(self._L.az_eval_rollout(self.h, C.c_uint64(seed), C.c_uint64(counter), self.stream), "eval_rollout")

    def search_rollout(self, iters, seed):
        """Whole vanilla search (ai/mcts.py:57-79): UCB1 + uniformly random rollouts on the device."""
        _lib.check(self._L.az_forest_search_rollout(self.h, int(iters), C.c_uint64(seed), self.stream), "search_rollout")

    def search_net(self, net, iters):
        """Whole search with the built-in net (ai.nn.B200Net): CUDA-graph replay per step."""
        _lib.check(self._L.az_forest_search_net(self.h, net.h, int(iters), self.stream), "search_net")

    def read_leaves(self):
        """Current evaluation batch and the evaluator's answers (host copies)."""
        m = self.n_trees * self.vl
        words = np.zeros((m, 2, self.g.nw), np.uint64)
        players = np.zeros(m, np.uint8)
        trees = np.zeros(m, np.int32)
        priors = np.zeros((m, self.g.action_dim), np.float32)
        values = np.zeros(m, np.float32)
        n = _lib.check(self._L.az_forest_read_leaves(self.h, m, _ptr(words), _ptr(players), _ptr(trees), _ptr(priors),
                                                     _ptr(values), self.stream), "read_leaves")
        return dict(n=n, words=words[:n], players=players[:n], trees=trees[:n], priors=priors[:n], values=values[:n])

    # ---- results --------------------------------------------------------------------------
    def root_visits(self):
        vis = np.zeros((self.n_trees, self.g.action_dim), dtype=np.int32)
        q = np.zeros(self.n_trees, dtype=np.float64)
        _lib.check(self._L.az_forest_root_visits(self.h, _ptr(vis), _ptr(q), self.stream), "root_visits")
        return vis, q

    def root_best(self):
        """Cell of root.max_visit_child() for every tree: first maximum in child order (ai/node.py:76-81)."""
        cells = np.zeros(self.n_trees, np.int32)
        _lib.check(self._L.az_forest_root_best(self.h, _ptr(cells), self.stream), "root_best")
        return cells

    def root_children(self, tree):
        m = self.g.max_children
        cells = np.zeros(m, np.int32)
        vis = np.zeros(m, np.int32)
        W = np.zeros(m, np.float64)
        P = np.zeros(m, np.float64)
        rv, rw = C.c_int(), C.c_double()
        n = _lib.check(self._L.az_forest_root_children(self.h, int(tree), m, _ptr(cells), _ptr(vis), _ptr(W), _ptr(P),
                                                       C.byref(rv), C.byref(rw), self.stream), "root_children")
        return dict(cells=cells[:n], visits=vis[:n], W=W[:n], P=P[:n], root_visit=rv.value, root_value=rw.value)

    # ---- evaluation cache (MCTS.cache, ai/mcts.py:7,37-51) ------------------------------------------
    def cache_config(self, log2_entries):
        """2**log2_entries entries; 0 turns the cache off (boards of <= 64 cells only)."""
        _lib.check(self._L.az_forest_cache_config(self.h, int(log2_entries)), "cache_config")

    def set_rounds(self, rounds):
        """vl == 1 with the cache: simulations a tree may run per launch while its leaves hit the cache."""
        _lib.check(self._L.az_forest_set_rounds(self.h, int(rounds)), "set_rounds")

    def cache_clear(self):
        _lib.check(self._L.az_forest_cache_clear(self.h, self.stream), "cache_clear")

    def cache_info(self):
        a, b = C.c_longlong(), C.c_longlong()
        _lib.check(self._L.az_forest_cache_info(self.h, C.byref(a), C.byref(b)), "cache_info")
        return dict(entries=a.value, bytes=b.value)

    def set_evaluator(self, ident):
        """Step-by-step searches: name the evaluator (any integer that changes with it) before begin_move."""
        _lib.check(self._L.az_forest_set_evaluator(self.h, C.c_ulonglong(int(ident) & (2 ** 64 - 1)), self.stream), "set_evaluator")

    def stats(self, reset=False):
        out = np.zeros(8, dtype=np.uint64)
        _lib.check(self._L.az_forest_stats(self.h, _ptr(out), int(reset), self.stream), "stats")
        keys = ["sims", "terminal", "expansions", "children", "selects", "child_reads", "collisions", "cache_hits"]
        return dict(zip(keys, out.tolist()))


def rules_legal(g, words):
    """Batched get_valid_moves on the GPU (differential tests).  words uint64 [n][2][nw]."""
    L = _lib.load()
    w = np.ascontiguousarray(words, dtype=np.uint64)
    n = w.shape[0]
    counts = np.zeros(n, np.int32)
    cells = np.zeros((n, g.state_dim), np.uint8)
    _lib.check(L.az_rules_legal(g.gid, g.rows, g.cols, n, _ptr(w), _ptr(counts), _ptr(cells)), "az_rules_legal")
    return [cells[i, :counts[i]].tolist() for i in range(n)]


def rules_winner(g, words, players, cells):
    L = _lib.load()
    w = np.ascontiguousarray(words, dtype=np.uint64)
    n = w.shape[0]
    pl = np.ascontiguousarray(players, np.int32)
    ce = np.ascontiguousarray(cells, np.int32)
    out = np.zeros(n, np.int32)
    _lib.check(L.az_rules_winner(g.gid, g.rows, g.cols, n, _ptr(w), _ptr(pl), _ptr(ce), _ptr(out)), "az_rules_winner")
    return out
