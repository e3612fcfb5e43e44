"""Experiment: split one GPU's concurrent games into P independent pipelines (forest + evaluator + stream each, one host
thread each) so that one pipeline's latency-bound advance kernel overlaps the other's tensor-core GEMMs.
    python tools/overlap_test.py [connect4|gomoku] [P ...]
Prints simulations/s (wall clock around all pipelines, synchronised on both sides) per P."""
import os
import sys
import threading
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
import alpha_zero_b200 as az  # noqa: E402
from alpha_zero_b200.ai.nn import Net, B200Net  # noqa: E402
from alpha_zero_b200.synth import random_roots  # noqa: E402


def run(wl, P, steps=4, warmup=2):
    G = bench.game_class(wl)
    n, iters, vl = wl["trees"] // P, wl["iters"], wl["vl"]
    A = G.action_dim
    NET = bench.NET
    torch.manual_seed(0)
    model = Net((G.rows, G.cols), NET["patch"], NET["dim"], A, NET["heads"], NET["depth"], G.feature_dim, dropout=0.1).eval()
    pipes = []
    rng = np.random.default_rng(7)
    for p in range(P):
        s = torch.cuda.Stream()
        eng = B200Net(model, max_batch=n * vl)
        forest = az.Forest(az.game_spec(wl["game"], wl["rows"]), n, iters, vl=vl, stream=s.cuda_stream)
        roots = [random_roots(G, n, wl["max_plies"], rng) for _ in range(steps + warmup)]
        noise = [rng.dirichlet([0.03] * A, size=n) for _ in range(steps + warmup)]
        pipes.append((s, eng, forest, roots, noise))

    def step(p, i):
        s, eng, forest, roots, noise = pipes[p]
        planes, players, mcs = roots[i]
        forest.set_roots_planes(planes, players, mcs)
        forest.set_noise(noise[i])
        forest.search_net(eng, iters)

    for i in range(warmup):                                   # sequential: graphs captured one pipeline at a time
        for p in range(P):
            step(p, i)
    torch.cuda.synchronize()

    def worker(p):
        for i in range(warmup, warmup + steps):
            step(p, i)

    t0 = time.perf_counter()
    th = [threading.Thread(target=worker, args=(p,)) for p in range(P)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    for p in range(P):
        v, _ = pipes[p][2].root_visits()
        assert int(v.sum()) == n * (iters - 1)
    return n * P * iters * steps / dt, dt / steps * 1e3


if __name__ == "__main__":
    name = sys.argv[1] if len(sys.argv) > 1 else "connect4"
    wl = bench.WORKLOADS[name]
    for P in [int(x) for x in sys.argv[2:]] or [1, 2]:
        v, ms = run(wl, P)
        print(f"{name} pipelines={P} games/pipeline={wl['trees'] // P}: {v / 1e6:.2f} M sims/s, {ms:.2f} ms per move", flush=True)
