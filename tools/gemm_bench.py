"""Warm-L2 timing of the cls-row GEMM shapes through az_gemm_bf16 (CUDA events, 200 launches each),
single-CTA tiles vs CTA-pair tiles (AZ_GEMM_BN = width, +1000 for the pair kernel).  GPU box only.
Epilogue ablation: build az_net.cu with -DAZ_EPI_EXPERIMENT into tools/_build/, point AZ_B200_LIB at it and set
AZ_GEMM_DBG bit 3 (no output stores) / bit 4 (no GELU); `python tools/gemm_bench.py <shape index> <bn>` runs one case."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import alpha_zero_b200 as az
L = az.lib.load()
shapes = [("out-proj", 4096, 512, 1024, 0), ("mlp1", 4096, 2048, 512, 1), ("mlp2", 4096, 512, 2048, 2), ("head-like", 4096, 64, 512, 0),
          ("mlp1 big", 65536, 2048, 512, 1), ("mlp2 big", 65536, 512, 2048, 2), ("mlp1 mid", 32768, 2048, 512, 1), ("out-proj mid", 32768, 512, 1024, 0),
          ("qkv full", 204800, 768, 256, 0), ("mlp1 full", 204800, 1024, 256, 1)]
import subprocess
if len(sys.argv) == 1:                                   # driver: one subprocess per (shape, tile) so a failing tile is isolated
    for i, (name, M, N, K, epi) in enumerate(shapes):
        for bn in (64, 128, 256, 1128, 1256):
            if N % (bn % 1000) == 0:
                r = subprocess.run([sys.executable, __file__, str(i), str(bn)], capture_output=True, text=True)
                print(r.stdout.strip() or f"{name} bn={bn}: FAILED rc={r.returncode} {r.stdout[-300:]} {r.stderr[-300:]}", flush=True)
    sys.exit(0)
if sys.argv[1] == "torch":                                # cuBLAS through torch.matmul: what a tuned library reaches on the same shapes
    for name, M, N, K, epi in shapes:
        a = torch.randn(M, K, device="cuda").to(torch.bfloat16)
        w = (torch.randn(N, K, device="cuda") / K ** 0.5).to(torch.bfloat16)
        for _ in range(10):
            torch.matmul(a, w.t())
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(200):
            torch.matmul(a, w.t())
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1000 / 200
        print(f"{name:10s} M{M} N{N} K{K} cuBLAS (no epilogue): {us:7.2f} us  {2.0 * M * N * K / us / 1e6:7.1f} TFLOP/s", flush=True)
    sys.exit(0)
shapes = [shapes[int(sys.argv[1])]]
bns = (int(sys.argv[2]),)
for name, M, N, K, epi in shapes:
    a = torch.randn(M, K, device="cuda").to(torch.bfloat16)
    w = (torch.randn(N, K, device="cuda") / K ** 0.5).to(torch.bfloat16)
    bias = torch.randn(N, device="cuda")
    resid = torch.randn(M, N, device="cuda").to(torch.bfloat16)
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    st = torch.cuda.current_stream().cuda_stream
    for bn in bns:
        w_ = bn % 1000
        if N % w_:
            continue
        os.environ["AZ_GEMM_BN"] = str(bn)
        def run():
            rc = L.az_gemm_bf16(a.data_ptr(), K, w.data_ptr(), bias.data_ptr(), resid.data_ptr(), N, out.data_ptr(), N, M, N, K, epi, st)
            assert rc == 0, rc
        for _ in range(10):
            run()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(200):
            run()
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1000 / 200
        print(f"{name:10s} M{M} N{N} K{K} bn={bn:5d} dbg={os.environ.get('AZ_GEMM_DBG', '0')}: {us:7.2f} us  {2.0 * M * N * K / us / 1e6:7.1f} TFLOP/s", flush=True)
