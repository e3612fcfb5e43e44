"""Vanilla mode (MCTS.mcts(None, ...): UCB1 + uniformly random rollouts, ai/mcts.py:57-79) on the device: rollout values of random
positions (a digest, to compare two builds of the library) and simulations/s of whole searches.
usage: [AZ_B200_LIB=other.so] python tools/rollout_bench.py"""
import hashlib, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import torch
import alpha_zero_b200 as az
import az_oracle as orc
from gpu_util import words_from_cells, random_positions

for g, n, plies, iters in ((orc.TICTACTOE, 4096, 3, 200), (orc.CONNECT4, 4096, 8, 400), (orc.GOMOKU, 4096, 8, 400)):
    rng = np.random.default_rng(3)
    cells, players, mcs = random_positions(g, n, rng, plies)
    f = az.Forest(az.game_spec(g.name, g.rows), n, iters, vl=1)
    f.set_roots_words(words_from_cells(g, cells), players, mcs)
    f.set_mode(True)
    f.begin_move(4)
    f.advance()
    f.eval_rollout(77, 0)
    vals = f.read_leaves()["values"].copy()
    f.set_roots_words(words_from_cells(g, cells), players, mcs)
    f.search_rollout(iters, 5)
    torch.cuda.synchronize()
    f.set_roots_words(words_from_cells(g, cells), players, mcs)
    t0 = time.time()
    f.search_rollout(iters, 5)
    torch.cuda.synchronize()
    dt = time.time() - t0
    vis, q = f.root_visits()
    print(f"{g.name:10s} rollout digest {hashlib.sha256(vals.tobytes()).hexdigest()[:16]} mean {vals.mean():+.4f} | search {n} trees x {iters} sims: "
          f"{n * iters / dt / 1e6:7.2f} M sims/s, visits digest {hashlib.sha256(vis.tobytes()).hexdigest()[:16]}")
    f.close()
