"""Per-kernel time of the cls-row evaluator at small batches: four separate GEMM launches (default), the chain kernel
(az_chain.cuh, AZ_CHAIN=1) and the feature-split cluster kernel (az_tail.cuh, AZ_FUSED_TAIL=1).  usage: python tools/tail_bench.py [connect4|gomoku] [boards ...]"""
import os, sys, subprocess
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

def run(game, sizes):
    import torch
    from alpha_zero_b200.ai.nn import Net, B200Net
    rows, cols, C, A = (6, 7, 3, 7) if game == "connect4" else (7, 7, 2, 49)
    torch.manual_seed(0)
    net = Net((rows, cols), 5, 512, A, 8, 1, C, dropout=0.1).eval()
    eng = B200Net(net, max_batch=4096)
    rng = np.random.default_rng(0)
    for B in sizes:
        x = (rng.random((B, C, rows, cols)) < 0.3).astype(np.float32)
        x[:, 1] *= (1 - x[:, 0])
        for _ in range(3):
            eng.forward_numpy(x)
        eng.profile(True)
        for _ in range(20):
            eng.forward_numpy(x)
        eng.profile(False)
        prof = eng.profile_read()
        parts = ", ".join(f"{k} {ms / cnt * 1e3:.1f} us x{cnt // 20}" for k, (ms, cnt) in prof.items() if cnt)
        print(f"flags {eng.path_flags():3d} {game} {B:5d} boards: {sum(ms for ms, _ in prof.values()) / 20 * 1e3:7.1f} us per pass | {parts}")
    eng.close()

if __name__ == "__main__":
    a = sys.argv[1:]
    if a and a[0] == "--child":
        run(a[1], [int(v) for v in a[2:]])
    else:
        game = a[0] if a else "connect4"
        sizes = a[1:] or ["256", "1024", "1792", "4096"]
        for env in ({}, {"AZ_CHAIN": "1"}, {"AZ_FUSED_TAIL": "1"}):
            print("==", env or "four launches (default)")
            sys.stdout.flush()
            subprocess.check_call([sys.executable, __file__, "--child", game] + sizes, env={**os.environ, **env})
