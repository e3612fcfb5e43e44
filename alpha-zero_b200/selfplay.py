"""Whole self-play games on the device, driven through the C ABI (csrc/az_selfplay.cu).

Host-side mirror of
  Game.self_play        games/gomoku.py:123-164 / connect4.py:117-151 / tictactoe.py:99-133
  train.collect_data    train.py:54-83 (the game loop)
with every per-ply decision (visit-count policy, np.random.choice / first max in child order, make_move,
check_winner, the next game in a finished slot, the Dirichlet root noise) taken by kernels: the host only
launches searches and reads a 16-byte status per ply.  Trajectories stay in HBM in compact form
(trajectory_arrays / device_tensors); games_from_arrays turns them into the reference's 5-tuples.
"""
import ctypes as C

import numpy as np

from . import lib as _lib
from .forest import Forest, game_spec


CONTINUOUS_ROUNDS = 4       # Connect4 4096 x 800 on B200: 46.1 / 59.4 / 65.8 / 67.2 / 68.3 / 66.1 / 63.6 M sims/s at 1 / 2 / 3 / 4 / 5 / 6 / 8


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def reference_sample_plies(Game, model=True):
    """Plies whose move is drawn from pi: Gomoku samples while move_count < 8 (gomoku.py:143), TicTacToe / Connect4
    sample every ply when a model is given and take the most visited child otherwise (connect4.py:133)."""
    if not model:
        return 0
    return 8 if Game.name == "gomoku" else Game.state_dim + 1


class SelfPlayEngine:
    """`n_games` games on the slots (trees) of `forest`; global id of local game k = id_first + k * id_stride."""

    def __init__(self, forest, n_games, seed, sample_plies, noise=True, id_first=0, id_stride=1):
        self.forest, self.n_games = forest, int(n_games)
        self.id_first, self.id_stride = int(id_first), int(id_stride)
        self._L = _lib.load()
        h = C.c_void_p()
        _lib.check(self._L.az_selfplay_create(forest.h, self.n_games, self.id_first, self.id_stride,
                                              C.c_uint64(int(seed) & (2 ** 64 - 1)), int(sample_plies), int(bool(noise)), C.byref(h)),
                   "az_selfplay_create")
        self.h = h
        self.g = forest.g
        self.max_plies = self.g.state_dim

    def close(self):
        if getattr(self, "h", None):
            self._L.az_selfplay_destroy(self.h)
            self.h = None

    __del__ = close

    # ---- driving ---------------------------------------------------------------------------------
    def begin(self, adopt_roots=False):
        _lib.check(self._L.az_selfplay_begin(self.h, int(bool(adopt_roots)), self.forest.stream), "az_selfplay_begin")

    def prepare(self):
        _lib.check(self._L.az_selfplay_prepare(self.h, self.forest.stream), "az_selfplay_prepare")

    def move(self):
        _lib.check(self._L.az_selfplay_move(self.h, self.forest.stream), "az_selfplay_move")

    def status(self):
        a, b, c, d = C.c_int(), C.c_int(), C.c_int(), C.c_longlong()
        _lib.check(self._L.az_selfplay_status(self.h, C.byref(a), C.byref(b), C.byref(c), C.byref(d), self.forest.stream),
                   "az_selfplay_status")
        return dict(started=a.value, finished=b.value, live=c.value, positions=d.value)

    def run(self, model, iters, max_plies=0, lockstep=False, max_steps=0, resume=False):
        """Plays until all games are over.  model: ai.nn.B200Net or ai.mcts.HashEvaluator.
        Default: continuous plies -- a slot whose search is finished moves and starts its next search at once, nobody waits
        for the slowest tree (returns the search steps run).  lockstep=True (or max_plies > 0: stop after that many
        searches) searches all slots together ply by ply (returns the searches run).  Both give the same trajectories."""
        from .ai.mcts import HashEvaluator
        from .ai.nn import B200Net
        if not lockstep and not max_plies:
            if isinstance(model, B200Net):
                return _lib.check(self._L.az_selfplay_run_async_net(self.h, model.h, int(iters), int(max_steps), int(bool(resume)),
                                                                    self.forest.stream), "az_selfplay_run_async_net")
            if isinstance(model, HashEvaluator):
                return _lib.check(self._L.az_selfplay_run_async_hash(self.h, int(iters), C.c_uint64(model.seed), int(model.qbits),
                                                                     self.forest.stream), "az_selfplay_run_async_hash")
        if isinstance(model, B200Net):
            return _lib.check(self._L.az_selfplay_run_net(self.h, model.h, int(iters), int(max_plies), self.forest.stream), "az_selfplay_run_net")
        if isinstance(model, HashEvaluator):
            return _lib.check(self._L.az_selfplay_run_hash(self.h, int(iters), C.c_uint64(model.seed), int(model.qbits), int(max_plies),
                                                           self.forest.stream), "az_selfplay_run_hash")
        raise TypeError("SelfPlayEngine.run needs a B200Net engine or a HashEvaluator")

    def launch_count(self):
        return int(self._L.az_selfplay_launch_count(self.h))

    # ---- results ----------------------------------------------------------------------------------
    def trajectory_arrays(self, with_uniforms=False):
        """Host copies: plies i32[G], winner i32[G], first i32[G], words u64[G,P,2,nw], visits u16[G,P,A], q f64[G,P],
        action (cell) i16[G,P] (and the move uniforms f64[G,P])."""
        G, P, g = self.n_games, self.max_plies, self.g
        out = dict(plies=np.zeros(G, np.int32), winner=np.zeros(G, np.int32), first=np.zeros(G, np.int32),
                   words=np.zeros((G, P, 2, g.nw), np.uint64), visits=np.zeros((G, P, g.action_dim), np.uint16),
                   q=np.zeros((G, P), np.float64), action=np.zeros((G, P), np.int16))
        u = np.zeros((G, P), np.float64) if with_uniforms else None
        _lib.check(self._L.az_selfplay_read(self.h, _ptr(out["plies"]), _ptr(out["winner"]), _ptr(out["first"]), _ptr(out["words"]),
                                            _ptr(out["visits"]), _ptr(out["q"]), _ptr(out["action"]), _ptr(u), self.forest.stream),
                   "az_selfplay_read")
        if with_uniforms:
            out["u"] = u
        return out

    def slots(self, with_noise=False):
        n, A = self.forest.n_trees, self.g.action_dim
        sg = np.zeros(n, np.int32)
        nz = np.zeros((n, A), np.float64) if with_noise else None
        _lib.check(self._L.az_selfplay_read_slots(self.h, _ptr(sg), _ptr(nz), self.forest.stream), "az_selfplay_read_slots")
        return (sg, nz) if with_noise else sg

    def device_tensors(self):
        """Zero-copy torch views of the trajectory store in HBM (for the NCCL all-gather): dict of CUDA tensors.
        u64 words are exposed as int64, u16 visits as int16 (bit patterns)."""
        import torch
        ptrs = (C.c_void_p * 8)()
        ng, mp = C.c_int(), C.c_int()
        _lib.check(self._L.az_selfplay_buffers(self.h, ptrs, C.byref(ng), C.byref(mp)), "az_selfplay_buffers")
        G, P, g = ng.value, mp.value, self.g

        class _View:
            def __init__(self, ptr, shape, typestr):
                self.__cuda_array_interface__ = dict(shape=shape, typestr=typestr, data=(int(ptr), False), version=2)

        def view(i, shape, typestr):
            return torch.as_tensor(_View(ptrs[i], shape, typestr), device=torch.device("cuda", torch.cuda.current_device()))
        return dict(plies=view(0, (G,), "<i4"), winner=view(1, (G,), "<i4"), words=view(2, (G, P, 2, g.nw), "<i8"),
                    visits=view(3, (G, P, g.action_dim), "<i2"), q=view(4, (G, P), "<f8"), action=view(5, (G, P), "<i2"),
                    u=view(6, (G, P), "<f8"), first=view(7, (G,), "<i4"), _owner=self)


# ---- compact arrays -> the reference's per-game tuples ------------------------------------------------
def words_to_planes(g, words, plies_parity=None):
    """u64 [..., 2, nw] -> float32 [..., C, rows, cols] boards as the reference holds them; for the 3-plane games
    plane 2 = side to move (connect4.py:55-70), given as plies_parity broadcastable to words.shape[:-2]."""
    w = np.asarray(words, dtype=np.uint64)
    lead = w.shape[:-2]
    bits = (w[..., None] >> np.arange(64, dtype=np.uint64)) & np.uint64(1)          # [..., 2, nw, 64]
    cells = bits.reshape(lead + (2, g.nw * 64))[..., :g.state_dim].astype(np.float32)
    out = np.zeros(lead + (g.planes, g.rows, g.cols), np.float32)
    out[..., 0, :, :] = cells[..., 0, :].reshape(lead + (g.rows, g.cols))
    out[..., 1, :, :] = cells[..., 1, :].reshape(lead + (g.rows, g.cols))
    if g.planes == 3:
        out[..., 2, :, :] = np.asarray(plies_parity, np.float32)[..., None, None]
    return out


def games_from_arrays(g, arr, return_visits=False):
    """Arrays of SelfPlayEngine.trajectory_arrays (or gathered ones) -> list of the reference's
    (boards, actions, policies, qs, winner) per game (gomoku.py:164), recorded plies only.  Vectorised over all
    positions; only the per-game slicing is a Python loop."""
    plies, first = np.asarray(arr["plies"]), np.asarray(arr["first"])
    G, P = arr["q"].shape
    parity = (np.arange(P) & 1)[None, :].repeat(G, 0)
    boards = words_to_planes(g, arr["words"], parity)
    vis = np.asarray(arr["visits"]).view(np.uint16).astype(np.int64)
    tot = vis.sum(axis=2, keepdims=True)
    pol = vis / np.where(tot == 0, 1, tot)                                             # utils.py:46-55 (float64)
    cell = np.asarray(arr["action"]).astype(np.int64)
    rr, cc = cell // g.cols, cell % g.cols
    games, visits = [], []
    for k in range(G):
        a, b = int(first[k]), int(plies[k])
        if b <= a and b == 0:
            games.append(([], [(-1, -1)], [], [], -2))
            visits.append(vis[k, :0])
            continue
        actions = [(-1, -1)] + list(zip(rr[k, a:b].tolist(), cc[k, a:b].tolist()))   # gomoku.py:128
        games.append((list(boards[k, a:b]), actions, list(pol[k, a:b]), arr["q"][k, a:b].tolist(), int(arr["winner"][k])))
        visits.append(vis[k, a:b])
    return (games, visits) if return_visits else games


def play_games_device(Game, model, n_games, mcts_iter, vl=1, dirichlet=True, concurrent=None, seed=None, id_first=0,
                      id_stride=1, return_arrays=False, return_engine=False, lockstep=False, rounds=None):
    """n_games self-play games on the device, `concurrent` in flight.  model: torch Net / B200Net / HashEvaluator.
    Continuous plies by default (lockstep=True: all slots search together ply by ply); rounds = cached simulations per tree
    and launch at vl == 1 (default: CONTINUOUS_ROUNDS for continuous plies, 1 in lockstep -- the measured optima)."""
    from .ai.mcts import MCTS, HashEvaluator
    slots = int(min(n_games, concurrent or n_games))
    forest = MCTS._forest_for(Game, slots, mcts_iter, vl)
    eng = model if isinstance(model, HashEvaluator) else MCTS._engine_for(model, forest.n_trees * forest.vl)
    if eng is None:
        raise TypeError("play_games_device needs a Net, a B200Net or a HashEvaluator")
    if seed is None:
        seed = int(np.random.randint(0, 2 ** 31 - 1)) * 2654435761 + int(np.random.randint(0, 2 ** 31 - 1))
    sp = SelfPlayEngine(forest, n_games, seed, reference_sample_plies(Game, True), noise=dirichlet, id_first=id_first, id_stride=id_stride)
    forest.set_active(None)
    forest.set_rounds(rounds if rounds else (CONTINUOUS_ROUNDS if (vl == 1 and not lockstep) else 1))
    sp.begin()
    try:
        sp.run(eng, mcts_iter, lockstep=lockstep)
    finally:
        forest.set_rounds(1)                                        # MCTS.mcts / mcts_batch on the same forest search in lockstep
    MCTS.mcts_count += mcts_iter * sp.status()["positions"]
    if return_engine:
        return sp
    arr = sp.trajectory_arrays()
    sp.close()
    return arr if return_arrays else games_from_arrays(forest.g, arr)
