"""The algebra behind csrc/az_front.cuh, checked in float64 on the CPU against the PyTorch Net.

For a depth-1 Net (ai/nn.py) the attention output of the cls token can be computed from ONE
contraction per token over the K = C*p*p + T binary inputs a_t = [im2col taps | one-hot(t)]:
    x_t - mean = Wc^T a_t                         (Wc = embedding rows centred over D)
    var_t      = a_t^T (Wc Wc^T) a_t / D          (Gram matrix; no cancellation, no mean needed)
    v_t        = r_t (Wv' Wc^T) a_t + bv'         (LayerNorm affine folded into Wv', bv')
    s_t,h      = r_t (Ws Wc^T)_h a_t              (constant cls query folded into Ws)
This test rebuilds those folded matrices exactly as az_net_finalize does and compares with the
module's own forward pass; it also bounds the error of rounding the folded weights to bf16."""
import numpy as np
import pytest
import torch


def _bf16(x):
    return torch.from_numpy(np.asarray(x, np.float32)).to(torch.bfloat16).to(torch.float32).numpy().astype(np.float64)


def _attention_out_reference(net, boards):
    """cls row of the attention sublayer's context (before out_proj), via torch in float64."""
    net = net.double()
    x = net.embedding(torch.from_numpy(boards).double())
    blk = net.blocks[0]
    h = blk.norm1(x)
    D = x.shape[-1]
    H = blk.attn.num_heads
    hd = D // H
    w, b = blk.attn.in_proj_weight, blk.attn.in_proj_bias
    q = (h[:, :1] @ w[:D].T + b[:D]).view(-1, 1, H, hd).transpose(1, 2)
    k = (h @ w[D:2 * D].T + b[D:2 * D]).view(h.shape[0], -1, H, hd).transpose(1, 2)
    v = (h @ w[2 * D:].T + b[2 * D:]).view(h.shape[0], -1, H, hd).transpose(1, 2)
    p = torch.softmax(q @ k.transpose(-1, -2) / hd ** 0.5, -1)
    return (p @ v).transpose(1, 2).reshape(h.shape[0], D).detach().numpy()


@pytest.mark.parametrize("rows,cols,C,A,D,P,H", [(6, 7, 3, 7, 128, 5, 8), (7, 7, 2, 49, 128, 5, 4), (3, 3, 3, 9, 128, 3, 8)])
def test_folded_front_end_equals_the_module(rows, cols, C, A, D, P, H):
    from alpha_zero_b200.ai.nn import Net
    torch.manual_seed(3)
    net = Net((rows, cols), P, D, A, H, 1, C, dropout=0.1).eval()
    with torch.no_grad():                                          # make the LayerNorm affine and the row means non-trivial
        net.blocks[0].norm1.weight.mul_(1.7)
        net.blocks[0].norm1.bias.add_(0.3)
        net.embedding.pos_embedding.add_(2.0)
    sd = {k: v.detach().double().numpy() for k, v in net.state_dict().items()}
    T, cells, kc, hd = rows * cols + 1, rows * cols, C * P * P, D // H
    B = 24
    rng = np.random.default_rng(1)
    boards = np.zeros((B, C, rows, cols), np.float32)
    for b in range(B):
        idx = rng.permutation(cells)[:rng.integers(0, cells)]
        for j, c in enumerate(idx):
            boards[b, j % 2, c // cols, c % cols] = 1
        if C == 3:
            boards[b, 2] = rng.integers(0, 2)
    with torch.no_grad():
        ref = _attention_out_reference(net, boards)
    # binary token rows a_t
    K = kc + T
    Amat = np.zeros((B, T, K))
    half = P // 2
    for t in range(1, T):
        r, c = (t - 1) // cols, (t - 1) % cols
        for ch in range(C):
            for kr in range(P):
                for kx in range(P):
                    rr, cc = r + kr - half, c + kx - half
                    if 0 <= rr < rows and 0 <= cc < cols:
                        Amat[:, t, ch * P * P + kr * P + kx] = boards[:, ch, rr, cc]
    for t in range(T):
        Amat[:, t, kc + t] = 1
    cw = sd["embedding.patch_embed.patch_embed.weight"].reshape(D, kc)
    cb, cls = sd["embedding.patch_embed.patch_embed.bias"], sd["embedding.cls_token"].reshape(D)
    pos = sd["embedding.pos_embedding"].reshape(T, D)
    We = np.zeros((K, D))
    We[:kc] = cw.T
    for t in range(T):
        We[kc + t] = pos[t] + (cls if t == 0 else cb)
    g1, be1 = sd["blocks.0.norm1.weight"], sd["blocks.0.norm1.bias"]
    wi, bi = sd["blocks.0.attn.in_proj_weight"], sd["blocks.0.attn.in_proj_bias"]
    wif, bif = wi * g1[None, :], bi + wi @ be1
    x0 = We[kc]
    q = wif[:D] @ ((x0 - x0.mean()) / np.sqrt(x0.var() + 1e-5)) + bif[:D]
    Ws = np.stack([(q[h * hd:(h + 1) * hd] @ wif[D + h * hd:D + (h + 1) * hd]) / np.sqrt(hd) for h in range(H)])
    Wc = We - We.mean(1, keepdims=True)
    G, W2v, W2s = Wc @ Wc.T, wif[2 * D:] @ Wc.T, Ws @ Wc.T

    def front(rnd):
        G_, Wv_, Ws_ = rnd(G), rnd(W2v), rnd(W2s)
        r = 1 / np.sqrt(np.maximum(np.einsum("btk,kj,btj->bt", Amat, G_, Amat), 0) / D + 1e-5)
        s = (Amat @ Ws_.T) * r[..., None]
        p = np.exp(s - s.max(1, keepdims=True))
        w = p / p.sum(1, keepdims=True) * r[..., None]
        z = Amat @ Wv_.T
        return np.concatenate([np.einsum("bt,btd->bd", w[..., h], z[..., h * hd:(h + 1) * hd]) for h in range(H)], 1) + bif[2 * D:]

    exact = front(lambda a: a)
    scale = np.abs(ref).max()
    assert np.abs(exact - ref).max() <= 1e-9 * max(scale, 1.0)          # identity in real arithmetic
    assert np.abs(front(_bf16) - ref).max() <= 2e-2 * max(scale, 1.0)   # folded weights rounded once to bf16


def test_layernorm_folded_into_the_consuming_gemm():
    """csrc/az_gemm.cuh: LayerNorm(x) W^T + b == rstd * (x W'^T - mean * colsum) + b' with W' = W diag(g),
    colsum[n] = sum_k W'[n][k], b' = b + W beta, mean / rstd from (sum, sum of squares) of the row.  Exact in
    real arithmetic; with W' rounded to bf16 and colsum formed from the ROUNDED weights the subtraction still
    cancels the mean term exactly (that is why az_net_finalize sums the rounded values)."""
    rng = np.random.default_rng(5)
    M, K, N = 64, 512, 96
    x = _bf16(rng.normal(0.4, 1.3, (M, K)))                          # rows as the producing GEMM stored them (bf16)
    W, b = rng.normal(0, 0.05, (N, K)), rng.normal(0, 0.1, N)
    g, beta = rng.normal(1.0, 0.2, K), rng.normal(0, 0.1, K)
    mean = x.mean(1, keepdims=True)
    var = x.var(1, keepdims=True)
    ref = ((x - mean) / np.sqrt(var + 1e-5) * g + beta) @ W.T + b

    s1, s2 = x.sum(1, keepdims=True), (x * x).sum(1, keepdims=True)
    m = s1 / K
    rstd = 1 / np.sqrt(np.maximum(s2 / K - m * m, 0) + 1e-5)
    Wf, bf = W * g[None, :], b + W @ beta
    out = rstd * (x @ Wf.T - m * Wf.sum(1)[None, :]) + bf
    assert np.abs(out - ref).max() <= 1e-9

    Wr = _bf16(Wf)                                                   # device weights; colsum from the rounded values
    out_r = rstd * (x @ Wr.T - m * Wr.sum(1)[None, :]) + bf
    assert np.abs(out_r - ref).max() <= 2e-2 * np.abs(ref).max()
    # colsum from UNROUNDED weights would leave mean * (rounding error of a whole row) behind: strictly worse
    out_bad = rstd * (x @ Wr.T - m * Wf.sum(1)[None, :]) + bf
    assert np.abs(out_r - ref).max() <= np.abs(out_bad - ref).max() + 1e-12


def test_polynomial_erf_gelu_is_far_below_bf16_resolution():
    """az_gelu_erf (csrc/az_gemm.cuh): gelu(x) = max(x, 0) - 0.5 |x| E(|x|/sqrt 2), E = 1 - erf from Abramowitz &
    Stegun 7.1.25.  Against the exact erf form of nn.GELU() the absolute error stays below 3e-5 everywhere, below half a
    bf16 ulp of the output wherever |gelu| >= 1e-2 (worst case: the shallow negative lobe near x = -2.5) and below
    1/16 ulp for |gelu| >= 0.25 -- the epilogue rounds to bf16 right after."""
    x = np.linspace(-9, 9, 200001)
    z = np.abs(x) / np.sqrt(2)
    t = 1 / (1 + 0.47047 * z)
    E = t * (0.3480242 + t * (-0.0958798 + t * 0.7478556)) * np.exp(-z * z)
    approx = np.maximum(x, 0) - 0.5 * np.abs(x) * E
    exact = torch.nn.functional.gelu(torch.from_numpy(x)).numpy()
    err = np.abs(approx - exact)
    assert err.max() <= 3e-5
    ulp = 2.0 ** (np.floor(np.log2(np.maximum(np.abs(exact), 2.0 ** -20))) - 7)
    big = np.abs(exact) >= 1e-2
    assert (err[big] <= ulp[big] / 2).all()
    large = np.abs(exact) >= 0.25
    assert (err[large] <= ulp[large] / 16).all()
