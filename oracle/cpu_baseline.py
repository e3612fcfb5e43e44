#!/usr/bin/env python
"""CPU baseline worker: the oracle port of the reference's search (oracle/az_oracle.py: one
Python thread, per-node objects, batch-1 torch forward per leaf -- the reference's design,
ai/mcts.py:11-60) timed on a bounded sample of the bench workload.  TEST/BENCH INFRASTRUCTURE.

Run as a subprocess by bench.py (`cpu_baseline` leg and `--impl reference`); prints one JSON
line {"sims": n, "seconds": s}.  Several workers run side by side, one per host core.
"""
import argparse
import json
import os
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(HERE))


def main():
    ap = argparse.ArgumentParser()
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
    import numpy as np
    import torch
    torch.set_num_threads(1)                       # the reference is single-threaded
    import az_oracle as orc
    from alpha_zero_b200.ai.nn import Net          # the reference-compatible torch module (CPU)

    g = orc.SPECS[a.game] if not a.rows or a.game != "gomoku" else orc.gomoku_spec(a.rows)
    torch.manual_seed(0)
    net = Net((g.rows, g.cols), a.patch, a.dim, g.action_dim, a.heads, a.depth, g.planes, dropout=0.1)
    net.eval()
    ev = orc.TorchEvaluator(net)
    rng = np.random.default_rng(a.seed)
    sims, t0 = 0, time.time()
    while time.time() - t0 < a.seconds:
        # a synthetic position like the GPU arm's: a few random legal plies, then ONE search,
        # cut into slices so the time budget is honoured
        board = orc.new_board(g)
        player, mc = 0, 0
        for _ in range(int(rng.integers(0, 8))):
            mv = orc.valid_moves(g, board)
            m = mv[int(rng.integers(len(mv)))]
            nxt = orc.make_move(g, board, player, m)
            if orc.check_winner(g, board, player, m) != -1:
                orc.undo_move(g, board, nxt, m)
                break
            player, mc = nxt, mc + 1
        root = orc.Node(None, None, player, mc)
        noise = rng.dirichlet([0.03] * g.action_dim)
        done = 0
        while done < a.iters and time.time() - t0 < a.seconds:
            step = min(25, a.iters - done)
            orc.mcts(g, ev, board, root, step, noise=noise)
            done += step
            sims += step
    print(json.dumps({"sims": sims, "seconds": time.time() - t0}))


if __name__ == "__main__":
    main()
