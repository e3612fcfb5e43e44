"""Phase clocks of az_advance_kernel over one whole search (library built with -DAZ_ADV_TRACE as
tools/libaz_trace.so by tools/build_trace.sh).  usage: python tools/adv_trace.py [connect4|gomoku] [trees] [iters] [vl]"""
import ctypes as C, sys, os
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from alpha_zero_b200 import lib
lib.SO_PATH = os.path.join(ROOT, "tools", "libaz_trace.so")
import alpha_zero_b200 as az
from alpha_zero_b200.ai.nn import Net, B200Net
from alpha_zero_b200.synth import random_roots
from alpha_zero_b200 import games

game = sys.argv[1] if len(sys.argv) > 1 else "connect4"
n = int(sys.argv[2]) if len(sys.argv) > 2 else (4096 if game == "connect4" else 8192)
iters = int(sys.argv[3]) if len(sys.argv) > 3 else 800
vl = int(sys.argv[4]) if len(sys.argv) > 4 else (1 if game == "connect4" else 8)
G = games.Connect4 if game == "connect4" else games.Gomoku
rows = 6 if game == "connect4" else 7
torch.manual_seed(0)
model = Net((G.rows, G.cols), 5, 512, G.action_dim, 8, 1, G.feature_dim, dropout=0.1).eval()
stream = torch.cuda.Stream()
eng = B200Net(model, max_batch=n * vl)
forest = az.Forest(az.game_spec(game, rows), n, iters, vl=vl, stream=stream.cuda_stream)
rng = np.random.default_rng(1)
L = lib.load()
L.az_debug_adv_trace.argtypes = [C.c_void_p, C.c_int]
buf = np.zeros((8192, 16), np.uint64)
names = ["tree scalars", "pend scalars+legal moves", "child block write", "backup", "root record", "select levels (sum)",
         "leaf hand-off", "terminal backup", "tail (slot atomic + writes)", "terminal test + path store + cache probe"]
for rep in range(2):
    planes, players, mcs = random_roots(G, n, 20 if game == "connect4" else 12, rng)
    forest.set_roots_planes(planes, players, mcs)
    forest.set_noise(rng.dirichlet([0.03] * G.action_dim, size=n))
    L.az_debug_adv_trace(buf.ctypes.data_as(C.c_void_p), 1)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(stream)
    forest.search_net(eng, iters)
    b.record(stream)
    torch.cuda.synchronize()
    L.az_debug_adv_trace(buf.ctypes.data_as(C.c_void_p), 0)
    t = buf[:n].astype(np.float64)
    launches = t[:, 15].mean()
    print(f"rep {rep}: search {a.elapsed_time(b):.1f} ms, launches/tree {launches:.0f}, levels/launch {t[:, 14].mean() / launches:.2f}")
    tot = 0.0
    for k, nm in enumerate(names):
        c = t[:, k].mean() / launches
        tot += c
        print(f"   {nm:28s} {c:9.0f} cycles/launch   (max tree {t[:, k].max() / launches:9.0f})")
    print(f"   {'sum':28s} {tot:9.0f} cycles/launch = {tot / 1.965e3:.1f} us;  per level {t[:, 5].mean() / max(t[:, 14].mean(), 1):.0f} cycles")

# per-launch distribution of warp lifetimes (eager steps)
L.az_debug_adv_hist.argtypes = [C.c_void_p]
hist = np.zeros(66, np.uint32)
planes, players, mcs = random_roots(G, n, 20 if game == "connect4" else 12, rng)
forest.set_roots_planes(planes, players, mcs)
forest.set_noise(rng.dirichlet([0.03] * G.action_dim, size=n))
forest.begin_move(iters)
steps = iters // vl + 1 if vl > 1 else iters
L.az_debug_adv_hist(hist.ctypes.data_as(C.c_void_p))
for s_ in range(steps):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(stream)
    forest.advance()
    b.record(stream)
    eng.eval_leaves(forest, stream.cuda_stream)
    if s_ % max(steps // 12, 1) == 0 or s_ == steps - 1:
        torch.cuda.synchronize()
        L.az_debug_adv_hist(hist.ctypes.data_as(C.c_void_p))
        h = hist[:64].astype(np.int64)
        tot = max(h.sum(), 1)
        cum = np.cumsum(h) / tot
        mean = (h * (np.arange(64) + 0.5) * 32768).sum() / tot
        q = lambda f: (int(np.searchsorted(cum, f)) + 1) * 32768 / 1965.0
        print(f"step {s_:4d}: kernel {a.elapsed_time(b) * 1e3:7.1f} us | warp life mean {mean / 1965:6.1f} us  p50 {q(0.5):6.1f}  p90 {q(0.9):6.1f}  p99 {q(0.99):6.1f}  max {hist[64] / 1965.0:7.1f} us  (n={tot})")
    elif s_ % max(steps // 12, 1) == max(steps // 12, 1) - 1:
        torch.cuda.synchronize()
        L.az_debug_adv_hist(hist.ctypes.data_as(C.c_void_p))      # reset so the next printed step is alone
