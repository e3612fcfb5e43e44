"""Max error of the B200 net against torch fp32, with and without the reassociated front kernel
(AZ_NO_FRONT2=1), and CUDA-event time of one forward pass.  GPU box only."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import torch
from alpha_zero_b200.ai.nn import B200Net
from test_gpu_net import make_net

CFG = [(6, 7, 3, 7, 512, 1, 5, 4096), (7, 7, 2, 49, 512, 1, 5, 8192), (6, 7, 3, 7, 256, 1, 5, 1111), (5, 5, 2, 25, 128, 1, 3, 777)]
for rows, cols, C, A, dim, depth, patch, B in CFG:
    net = make_net(rows, cols, C, A, dim, depth, patch)
    rng = np.random.default_rng(0)
    x = (rng.random((B, C, rows, cols)) < 0.3).astype(np.float32)
    x[:, 1] *= (1 - x[:, 0])
    if C == 3:
        x[:, 2] = rng.integers(0, 2, B).astype(np.float32)[:, None, None]
    x[0] = 0
    with torch.no_grad():
        rl, rv = net(torch.from_numpy(x[:600]))
    rl, rv = rl.numpy(), rv.numpy()
    for mode in ("front2", "front1"):
        if mode == "front1":
            os.environ["AZ_NO_FRONT2"] = "1"
        else:
            os.environ.pop("AZ_NO_FRONT2", None)
        eng = B200Net(net, max_batch=B)
        l, v, p = eng.forward_numpy(x)
        t0 = time.perf_counter()
        for _ in range(5):
            eng.forward_numpy(x)
        dt = (time.perf_counter() - t0) / 5
        el = np.abs(l[:600] - rl)
        print(f"{rows}x{cols} D{dim} B{B} {mode}: max|dlogit| {el.max():.4f} (tol-slack {(0.05 + 0.02 * np.abs(rl) - el).min():.4f}) "
              f"max|dvalue| {np.abs(v[:600] - rv).max():.4f} nan {int(np.isnan(l).sum())} fwd_numpy {dt * 1e3:.2f} ms", flush=True)
        eng.close()
