"""tcgen05 GEMM (csrc/az_gemm.cuh) through the C ABI vs a plain PyTorch fp32 matmul of the
same bf16-rounded operands.  Tolerance: the output is rounded to bf16 (2^-8 relative) on top
of fp32 accumulation-order differences -> |err| <= 1e-2 * (|ref| + 1)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def run_gemm(M, N, K, epi, seed=0, lda=None, ldo=None):
    import torch
    import alpha_zero_b200 as az
    L = az.lib.load()
    g = torch.Generator(device="cuda").manual_seed(seed)
    lda = lda or K
    ldo = ldo or N
    a_full = torch.randn(M, lda, device="cuda", generator=g).to(torch.bfloat16)
    w = (torch.randn(N, K, device="cuda", generator=g) / K ** 0.5).to(torch.bfloat16)
    bias = torch.randn(N, device="cuda", generator=g)
    resid = torch.randn(M, N, device="cuda", generator=g).to(torch.bfloat16)
    out = torch.full((M, ldo), 7.0, device="cuda", dtype=torch.bfloat16)
    rc = L.az_gemm_bf16(a_full.data_ptr(), lda, w.data_ptr(), bias.data_ptr(), resid.data_ptr(), N, out.data_ptr(),
                        ldo, M, N, K, epi, torch.cuda.current_stream().cuda_stream)
    az.lib.check(rc, "az_gemm_bf16")
    torch.cuda.synchronize()
    ref = a_full[:, :K].float() @ w.float().t() + bias
    if epi == 1:
        ref = torch.nn.functional.gelu(ref)
    if epi == 2:
        ref = ref + resid.float()
    got = out[:, :N].float()
    err = (got - ref).abs()
    tol = 1e-2 * (ref.abs() + 1.0)
    bad = (err > tol).sum().item()
    assert bad == 0, f"{bad} / {M * N} elements off; max err {err.max().item()}"
    if ldo > N:
        assert (out[:, N:] == 7.0).all()


@pytest.mark.parametrize("M,N,K,epi", [
    (128, 64, 64, 0), (128, 128, 64, 0), (128, 256, 64, 0),      # one tile, one k-block
    (128, 256, 512, 0), (256, 512, 512, 0),                       # k loop, several tiles
    (300, 192, 128, 0), (77, 64, 192, 1), (1000, 1536, 512, 2),    # ragged M, all epilogues
    (4096 * 43, 1024, 512, 0),                                      # Connect4 KV projection at full size
    (8192, 2048, 512, 1), (8192, 512, 2048, 2),                     # MLP shapes
])
def test_gemm_matches_torch(M, N, K, epi):
    run_gemm(M, N, K, epi)


def test_gemm_strided_rows():
    run_gemm(512, 512, 512, 0, lda=512 * 3, ldo=1024)
