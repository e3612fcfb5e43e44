"""Per-kernel time of the generic network path (full blocks before the cls-row tail) with the tcgen05 all-token attention
and, under AZ_NO_ATTN_MMA=1, with the scalar kernel it replaces.  usage: python tools/attn_bench.py [rows] [dim] [depth] [patch] [boards]"""
import os, sys, subprocess
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

def run(rows, dim, depth, patch, B):
    import torch
    from alpha_zero_b200.ai.nn import Net, B200Net
    torch.manual_seed(0)
    net = Net((rows, rows), patch, dim, rows * rows, 8, depth, 2, dropout=0.1).eval()
    eng = B200Net(net, max_batch=B)
    rng = np.random.default_rng(0)
    x = (rng.random((B, 2, rows, rows)) < 0.3).astype(np.float32)
    x[:, 1] *= (1 - x[:, 0])
    for _ in range(3):
        eng.forward_numpy(x)
    eng.profile(True)
    for _ in range(10):
        eng.forward_numpy(x)
    eng.profile(False)
    prof = eng.profile_read()
    tot = sum(ms for ms, _ in prof.values())
    print(f"flags {eng.path_flags()}  rows {rows} D {dim} depth {depth} boards {B}: {tot / 10:.3f} ms per forward")
    for k, (ms, cnt) in sorted(prof.items(), key=lambda kv: -kv[1][0]):
        if cnt:
            print(f"   {k:12s} {ms / 10:8.3f} ms/forward  {cnt // 10:3d} launches  {ms / cnt * 1e3:8.1f} us each")
    eng.close()

if __name__ == "__main__":
    a = sys.argv[1:]
    if a and a[0] == "--child":
        run(*[int(v) for v in a[1:]])
    else:
        args = [str(v) for v in ([int(v) for v in a] + [7, 256, 2, 5, 4096][len(a):])]
        for env in ({}, {"AZ_NO_ATTN_MMA": "1"}):
            print("==", env or "tcgen05 attention")
            sys.stdout.flush()
            subprocess.check_call([sys.executable, __file__, "--child"] + args, env={**os.environ, **env})
