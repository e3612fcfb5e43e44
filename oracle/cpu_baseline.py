#!/usr/bin/env python
"""CPU baseline worker (TEST/BENCH INFRASTRUCTURE): the reference's self-play search timed on the host cores.

Two kinds, one JSON line {"sims", "seconds", "kind", "games", "cache_hit_rate"} each:

  --kind reference   the UNMODIFIED reference from baseline/_ref (a git-ignored copy of /root/reference made by
                     __graft_entry__.build(); it travels to the GPU box with the snapshot): Gomoku().self_play(model,
                     iters) with the reference's own Net (ai/nn.py), MCTS.cache on, simulations counted with the
                     reference's own counter MCTS.mcts_count (ai/mcts.py:9,17).  Only Gomoku runs unshimmed with a
                     network, so this kind is Gomoku only.
  --kind port        the oracle port (oracle/az_oracle.py: one Python thread, per-node objects, batch-1 torch forward
                     per leaf, an evaluation cache like MCTS.cache) playing whole self-play games -- used where the
                     reference cannot run (Connect4 / TicTacToe with a network) or baseline/_ref is absent.

Whole games are played back to back (so the cache sees the cross-move reuse the reference gets) until --seconds have
passed; the run is cut by a timer wherever it is, the simulations completed so far count.
"""
import argparse
import json
import os
import signal
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = os.path.join(ROOT, "baseline", "_ref")


class _Timeout(Exception):
    pass


def _alarm(_sig, _frm):
    raise _Timeout()


def run_reference(a):
    """The reference's own code path, nothing of this repository on it."""
    sys.path.insert(0, REF)                          # its modules are top-level (`utils`, `ai`, `games`, ...)
    os.chdir(REF)                                    # importing `games` creates logs/ next to it (utils.py:71-90)
    import numpy as np
    import torch
    torch.set_num_threads(1)                         # the reference is single-threaded
    from ai import MCTS                              # noqa: the reference's classes
    from ai.nn import Net
    from games import Gomoku
    np.random.seed(a.seed)
    torch.manual_seed(0)
    model = Net(Gomoku.rows, a.patch, a.dim, Gomoku.action_dim, a.heads, a.depth, Gomoku.feature_dim, 0.1)   # main.py:134
    model.eval()
    games = 0
    signal.signal(signal.SIGALRM, _alarm)
    t0 = time.time()
    signal.setitimer(signal.ITIMER_REAL, a.seconds)
    try:
        with torch.no_grad():                        # train.py:63
            while True:
                Gomoku().self_play(model, a.iters)
                games += 1
    except _Timeout:
        pass
    dt = time.time() - t0
    sims = int(MCTS.mcts_count)
    evals = max(sims, 1)
    return dict(sims=sims, seconds=dt, kind="reference", games=games, cache_hit_rate=float(MCTS.matched) / evals,
                cache_entries=len(MCTS.cache))


def run_port(a):
    sys.path.insert(0, HERE)
    sys.path.insert(0, ROOT)
    import numpy as np
    import torch
    torch.set_num_threads(1)
    import az_oracle as orc
    from alpha_zero_b200.ai.nn import Net            # the reference-compatible torch module (CPU, fp32)

    g = orc.SPECS[a.game] if not a.rows or a.game != "gomoku" else orc.gomoku_spec(a.rows)
    torch.manual_seed(0)
    net = Net((g.rows, g.cols), a.patch, a.dim, g.action_dim, a.heads, a.depth, g.planes, dropout=0.1)
    net.eval()
    inner = orc.TorchEvaluator(net)
    cache, stat = {}, dict(hit=0, evals=0, sims=0)

    def evaluator(canon):                            # MCTS.cache (ai/mcts.py:37-51): canonical_board.tobytes() -> answer
        key = canon.tobytes()
        stat["evals"] += 1
        got = cache.get(key)
        if got is None:
            got = inner(canon)
            cache[key] = got
        else:
            stat["hit"] += 1
        return got

    real_backup = orc.backup

    def counting_backup(*args, **kw):                # one backup per simulation (ai/mcts.py:29,32,60)
        stat["sims"] += 1
        return real_backup(*args, **kw)
    orc.backup = counting_backup
    rng = np.random.RandomState(a.seed)
    games = 0
    signal.signal(signal.SIGALRM, _alarm)
    t0 = time.time()
    signal.setitimer(signal.ITIMER_REAL, a.seconds)
    try:
        while True:
            orc.self_play(g, evaluator, a.iters, noise_fn=lambda A: rng.dirichlet([0.03] * A), rng=rng)
            games += 1
    except _Timeout:
        pass
    dt = time.time() - t0
    return dict(sims=stat["sims"], seconds=dt, kind="port", games=games, cache_hit_rate=stat["hit"] / max(stat["evals"], 1),
                cache_entries=len(cache))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--kind", default="port", choices=["port", "reference"])
    ap.add_argument("--game", default="connect4")
    ap.add_argument("--rows", type=int, default=0)
    ap.add_argument("--dim", type=int, default=512)
    ap.add_argument("--depth", type=int, default=1)
    ap.add_argument("--patch", type=int, default=5)
    ap.add_argument("--heads", type=int, default=8)
    ap.add_argument("--iters", type=int, default=800)
    ap.add_argument("--seconds", type=float, default=15.0)
    ap.add_argument("--seed", type=int, default=0)
    a = ap.parse_args()
    os.environ.setdefault("OMP_NUM_THREADS", "1")
    if a.kind == "reference":
        if a.game != "gomoku" or (a.rows and a.rows != 7):
            raise SystemExit("--kind reference: only Gomoku 7x7 runs unshimmed with a network")
        if not os.path.isdir(os.path.join(REF, "games")):
            raise SystemExit("baseline/_ref is missing (run `python __graft_entry__.py build` where /root/reference exists)")
        out = run_reference(a)
    else:
        out = run_port(a)
    sys.stdout.write(json.dumps(out) + "\n")


if __name__ == "__main__":
    main()
