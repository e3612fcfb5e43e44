"""MCTS.mcts with the reference's signature (ai/mcts.py:11), executed on the GPU.

    MCTS.mcts(model, board, root, Game, mcts_iterations, dirichlet=True)

`board` (float32 [C,H,W]) is left untouched (the reference mutates and restores it); results
are read from `root` afterwards: root.visit / root.value / root.children[i].{visit, value, prior,
prevAction}, in the reference's child order.  `model` may be
  * a B200Net engine or a torch `Net` (uploaded once, re-uploaded when its weights change):
    the whole search is one CUDA-graph replay loop inside libaz_b200.so;
  * any other callable x[B,C,H,W] -> (logits[B,A], values[B,1]): leaves are handed to it as a
    float32 CUDA tensor batch each step (external evaluator);
  * a falsy value selects the reference's vanilla mode (ai/mcts.py:57-79): UCB1 selection (bit-exact) and
    uniformly random rollouts on the device (az_forest_search_rollout; statistical parity).
MCTS.mcts_batch does the same for many independent positions at once (the throughput path).
"""
import numpy as np

from ..forest import Forest, game_spec
from .node import Node


def _spec_of(Game):
    return game_spec(Game.name, Game.rows)


class HashEvaluator:
    """Build-defined deterministic evaluator (exact in fp32, shared with the oracle) used for
    parity runs: MCTS.mcts(HashEvaluator(seed), ...) searches with az_eval_hash on the device."""

    def __init__(self, seed=0, qbits=16):
        self.seed, self.qbits = int(seed), int(qbits)


class UCB1HashEvaluator(HashEvaluator):
    """Parity runs of the vanilla selection rule (model=None, utils.py:36,43): UCB1 on the device with the
    deterministic hash value in place of the random rollout (oracle: az_oracle.mcts_vanilla + hash_simulate)."""


class MCTS:
    cache = {}          # kept for API compatibility (main.py:51-57); the GPU path needs no cache
    matched = 0
    mcts_count = 0
    _forests = {}
    _engines = {}

    # ---- evaluator plumbing -----------------------------------------------------------------
    @staticmethod
    def _engine_for(model, max_batch):
        from .nn import Net, B200Net
        if isinstance(model, B200Net):
            return model
        if isinstance(model, Net):
            import weakref
            stamp = tuple(p._version for p in model.parameters())
            ent = MCTS._engines.get(id(model))
            if ent is not None and ent[3]() is not model:   # id() of a collected Net reused by another model
                ent[0].close()
                ent = None
            if ent is not None and ent[2] < max_batch:      # a larger forest than the engine was sized for
                ent[0].close()
                ent = None
            if ent is None:
                for k in [k for k, e in MCTS._engines.items() if e[3]() is None]:   # engines of collected models
                    MCTS._engines.pop(k)[0].close()
                ent = [B200Net(model, max_batch=max_batch), stamp, max_batch, weakref.ref(model)]
                MCTS._engines[id(model)] = ent
            elif ent[1] != stamp:                       # weights changed since the upload
                ent[0].load_state_dict(model.state_dict())
                ent[1] = stamp
            return ent[0]
        return None

    _MAX_FORESTS = 3            # node pools are large (Gomoku 8192 x 800: 10 GB): keep the most recently used few

    @staticmethod
    def _forest_for(Game, n_trees, iters, vl, tag=0):
        spec = _spec_of(Game)
        key = (spec.name, spec.rows, n_trees, vl, tag)
        f = MCTS._forests.pop(key, None)
        if f is None or f.max_iters < iters:
            if f is not None:
                f.close()
            while len(MCTS._forests) >= MCTS._MAX_FORESTS:            # least recently used first (dicts keep insertion order)
                MCTS._forests.pop(next(iter(MCTS._forests))).close()
            f = Forest(spec, n_trees, max(iters, 1), vl=vl)
        MCTS._forests[key] = f                                        # (re)inserted last = most recently used
        return f

    @staticmethod
    def release():
        """Frees every cached forest and inference engine (device memory); they are rebuilt on demand."""
        for f in MCTS._forests.values():
            f.close()
        MCTS._forests.clear()
        for ent in MCTS._engines.values():
            ent[0].close()
        MCTS._engines.clear()

    @staticmethod
    def _search_external(forest, model, iters):
        """Arbitrary callable evaluator: one Python round trip per step."""
        import torch
        g = forest.g
        m = forest.n_trees * forest.vl
        planes = torch.zeros((m, g.planes, g.rows, g.cols), dtype=torch.float32, device="cuda")
        dev = next((p.device for p in getattr(model, "parameters", lambda: [])()), torch.device("cpu"))
        L = forest._L
        from .. import lib as _lib
        stamp = hash((id(model),) + tuple(p._version for p in getattr(model, "parameters", lambda: [])())) & (2 ** 63 - 1)
        _lib.check(L.az_forest_set_evaluator(forest.h, stamp | (1 << 62), forest.stream), "set_evaluator")
        forest.begin_move(iters)
        for _ in range(2 * iters + 8):
            forest.advance()
            n_leaves, n_done = forest.counts()
            if n_leaves:
                _lib.check(L.az_leaves_to_planes(forest.h, planes.data_ptr(), m, forest.stream), "leaves_to_planes")
                torch.cuda.synchronize()
                with torch.no_grad():
                    logits, values = model(planes[:n_leaves].to(dev))
                logits = logits.to("cuda", torch.float32).contiguous()
                values = values.to("cuda", torch.float32).reshape(-1).contiguous()
                torch.cuda.synchronize()
                _lib.check(L.az_leaves_set_logits(forest.h, logits.data_ptr(), values.data_ptr(), n_leaves, forest.stream),
                           "leaves_set_logits")
            if n_done >= forest.n_trees and n_leaves == 0:
                return
        raise RuntimeError("external-evaluator search did not converge")

    @staticmethod
    def _run(forest, model, iters):
        if not model:
            # vanilla MCTS (ai/mcts.py:57-79): UCB1 + random rollouts on the device; the rollout streams are seeded
            # from the global NumPy RNG the reference draws its rollout moves from
            forest.set_mode(True)
            try:
                forest.search_rollout(iters, int(np.random.randint(0, 2 ** 31 - 1)) * 2654435761 + 12345)
            finally:
                forest.set_mode(False)
            return
        if isinstance(model, UCB1HashEvaluator):               # parity runs of the vanilla selection rule
            forest.set_mode(True)
            try:
                forest.search_hash(iters, model.seed, model.qbits)
            finally:
                forest.set_mode(False)
            return
        if isinstance(model, HashEvaluator):
            forest.search_hash(iters, model.seed, model.qbits)
            return
        eng = MCTS._engine_for(model, forest.n_trees * forest.vl)
        if eng is not None:
            forest.search_net(eng, iters)
        else:
            MCTS._search_external(forest, model, iters)

    # ---- the reference entry point ---------------------------------------------------------------
    @staticmethod
    def mcts(model, board, root, Game, mcts_iterations, dirichlet=True):
        forest = MCTS._forest_for(Game, 1, mcts_iterations, 1)
        forest.set_roots_planes(np.asarray(board, dtype=np.float32)[None], [root.currentPlayer], [root.move_count])
        # the same draw, from the same global RNG, as utils.add_dirichlet_noise (utils.py:21)
        # (vanilla mode, model=None, never mixes noise in: ai/mcts.py:42,52 sit inside `if model`)
        forest.set_noise(np.random.dirichlet([0.03] * Game.action_dim)[None] if (dirichlet and model) else None)
        MCTS._run(forest, model, mcts_iterations)
        MCTS.mcts_count += mcts_iterations
        ch = forest.root_children(0)
        root.visit, root.value = ch["root_visit"], ch["root_value"]
        root.children = []
        for cell, n, w, p in zip(ch["cells"], ch["visits"], ch["W"], ch["P"]):
            action = (int(cell) // Game.cols, int(cell) % Game.cols)
            node = Node(root, action, 1 - root.currentPlayer, root.move_count + 1,
                        (np.float64(p) if dirichlet else np.float32(p)) if model else 0.0)
            node.visit, node.value = int(n), float(w)
            root.children.append(node)
        root.refresh_ucb("network" if model else "normal")

    # ---- many positions at once --------------------------------------------------------------------
    @staticmethod
    def mcts_batch(model, boards, players, move_counts, Game, mcts_iterations, noise=None, vl=1, active=None, forest_tag=0,
                   return_best=False):
        """boards float32 [n,C,H,W]; noise float64 [n,A] or None.  Returns (visits int32 [n,A],
        q float64 [n]) -- the visit-count policy numerators and root.value/root.visit -- and with return_best the cell of
        root.max_visit_child() per tree (first maximum in child order).  forest_tag selects a separate forest (and
        evaluation cache) for callers that alternate between evaluators, e.g. the arena."""
        boards = np.asarray(boards, dtype=np.float32)
        forest = MCTS._forest_for(Game, boards.shape[0], mcts_iterations, vl, forest_tag)
        forest.set_roots_planes(boards, players, move_counts)
        forest.set_noise(noise)
        forest.set_active(active)
        MCTS._run(forest, model, mcts_iterations)
        MCTS.mcts_count += mcts_iterations * boards.shape[0]
        if return_best:
            return forest.root_visits() + (forest.root_best(),)
        return forest.root_visits()
