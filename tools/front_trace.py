"""Timeline of one CTA of az_front_kernel (library built with -DAZ_FRONT_TRACE)."""
import ctypes as C, sys, os
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from alpha_zero_b200.ai.nn import Net, B200Net
from alpha_zero_b200 import lib
rows, cols, Cc, A = (6, 7, 3, 7) if len(sys.argv) < 2 or sys.argv[1] == "c4" else (7, 7, 2, 49)
torch.manual_seed(0)
net = Net((rows, cols), 5, 512, A, 8, 1, Cc, dropout=0.1).eval()
B = 4096
eng = B200Net(net, max_batch=B)
rng = np.random.default_rng(0)
x = (rng.random((B, Cc, rows, cols)) < 0.3).astype(np.float32)
x[:, 1] *= (1 - x[:, 0])
for _ in range(3):
    eng.forward_numpy(x)
L = lib.load()
out = np.zeros((4, 64, 8), np.uint64)
L.az_debug_front_trace.argtypes = [C.c_void_p]
print("rc", L.az_debug_front_trace(out.ctypes.data_as(C.c_void_p)))
t0 = out[out > 0].min()
names = ["mma", "builder", "S", "V|U"]
for it in range(0, 14):
    for r in range(4):
        ev = [int(v - t0) if v else -1 for v in out[r, it]]
        print(f"it {it:2d} {names[r]:8s}", ev[:8])
