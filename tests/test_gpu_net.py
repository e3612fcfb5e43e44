"""The B200 policy/value net (tcgen05 GEMMs + fused kernels) against the reference's Net.

Floating point: operands are bf16 (2^-8 relative rounding) with fp32 accumulation, through at
most `depth` blocks, so logits are compared with |err| <= 0.05 + 0.02*|ref|, tanh values with
|err| <= 0.03 and softmax priors with |err| <= 5e-3 (measured: 0.022 / 0.015 / 1e-3).
Search parity WITH the net is exact: the GPU's own priors/values are replayed in the C oracle
through the evaluator table (the reference's MCTS.cache preload route, ai/mcts.py:40-41).
"""
import json
import os

import numpy as np
import pytest

import az_oracle as orc
import c_oracle as co
from golden_util import GOLDEN
from gpu_util import words_from_cells, random_positions, compare_forest_with_oracle

pytestmark = pytest.mark.gpu


def make_net(rows, cols, C, A, dim, depth, patch, heads=8, seed=5):
    import torch
    from alpha_zero_b200.ai.nn import Net
    from gen_golden import deterministic_state_dict
    net = Net((rows, cols), patch, dim, A, heads, depth, C, dropout=0.1)
    net.load_state_dict(deterministic_state_dict(net, seed))
    net.eval()
    return net


def check_close(l, v, p, rl, rv):
    assert (np.abs(l - rl) <= 0.05 + 0.02 * np.abs(rl)).all(), np.abs(l - rl).max()
    assert (np.abs(v - rv) <= 0.03).all(), np.abs(v - rv).max()
    rp = np.exp(rl) / np.exp(rl).sum(1, keepdims=True)          # ai/mcts.py:49
    assert (np.abs(p - rp) <= 5e-3).all(), np.abs(p - rp).max()
    assert np.allclose(p.sum(1), 1.0, atol=1e-5)


def test_net_matches_reference_goldens():
    """Outputs of the unmodified reference Net (tests/golden/net.npz) on weights that are a
    pure function of (name, shape, seed)."""
    from alpha_zero_b200.ai.nn import Net, B200Net
    from gen_golden import deterministic_state_dict
    recs = json.load(open(os.path.join(GOLDEN, "net.json")))
    arr = np.load(os.path.join(GOLDEN, "net.npz"))
    for i, c in enumerate(recs):
        net = Net(c["img"], c["patch"], c["dim"], c["A"], c["heads"], c["depth"], c["C"], dropout=0.1)
        assert sorted(net.state_dict().keys()) == c["keys"]
        net.load_state_dict(deterministic_state_dict(net, c["wseed"]))
        eng = B200Net(net, max_batch=64)
        l, v, p = eng.forward_numpy(arr[f"x{i}"])
        check_close(l, v, p, arr[f"logits{i}"], arr[f"values{i}"])
        eng.close()


@pytest.mark.parametrize("rows,cols,C,A,dim,depth,patch,B", [
    (6, 7, 3, 7, 512, 1, 5, 5000),        # Connect4, the main.py:134 net shape on a 6x7 board
    (7, 7, 2, 49, 512, 1, 5, 9000),       # Gomoku, main.py:134
    (7, 7, 2, 49, 256, 2, 5, 1500),       # main.py:187-188 (depth 2: full block + pruned block)
    (3, 3, 3, 9, 64, 1, 3, 700),          # TicTacToe small net
    (3, 3, 3, 9, 128, 1, 3, 700),         # TicTacToe on the fused front end (8 boards per tile)
    (6, 7, 3, 7, 256, 1, 5, 1111),        # Connect4, two feature blocks, ragged last tile
    (15, 15, 2, 225, 128, 2, 3, 300),     # 15x15: T = 226 tokens (head dim 16, four key boxes, two query tiles)
    (7, 7, 2, 49, 512, 2, 5, 700),        # depth 2 at D = 512: head dim 64, one head per column group
    (9, 9, 2, 81, 256, 3, 3, 500),        # 9x9, depth 3: T = 82 tokens (two key boxes), two full blocks
])
def test_net_matches_torch_fp32(rows, cols, C, A, dim, depth, patch, B):
    import torch
    from alpha_zero_b200.ai.nn import B200Net
    net = make_net(rows, cols, C, A, dim, depth, patch)
    rng = np.random.default_rng(0)
    x = (rng.random((B, C, rows, cols)) < 0.3).astype(np.float32)
    x[:, 1] *= (1 - x[:, 0])
    if C == 3:
        x[:, 2] = rng.integers(0, 2, B).astype(np.float32)[:, None, None]
    x[0] = 0                                                   # empty board
    eng = B200Net(net, max_batch=4096)
    assert bool(eng.path_flags() & 32) == (depth > 1), "full blocks must run their attention on the tensor cores"
    l, v, p = eng.forward_numpy(x)
    with torch.no_grad():
        rl, rv = net(torch.from_numpy(x[:600]))
    check_close(l[:600], v[:600], p[:600], rl.numpy(), rv.numpy())
    l2, v2, p2 = eng.forward_numpy(x[-100:])                     # position in the batch is irrelevant
    assert np.array_equal(l2, l[-100:]) and np.array_equal(v2, v[-100:])
    lt, vt = eng(torch.from_numpy(x[:3]))                        # the `model(...)` calling convention
    assert lt.shape == (3, A) and vt.shape == (3, 1)
    eng.close()


@pytest.mark.parametrize("env", [{"AZ_PAIR_MIN_ROWS": "512"}, {"AZ_NO_FRONT2": "1"}, {"AZ_NO_FRONT2": "1", "AZ_PAIR_MIN_ROWS": "512"}])
def test_net_kernel_variants_match_torch(env, monkeypatch):
    """The same tolerance on the other kernel choices of the cls-row path: CTA-pair GEMM tiles (az_gemm2.cuh, default
    only from 16384 boards per pass) and the first fused front kernel (az_front.cuh) instead of the reassociated one."""
    import torch
    from alpha_zero_b200.ai.nn import B200Net
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    net = make_net(6, 7, 3, 7, 512, 1, 5)
    rng = np.random.default_rng(2)
    B = 3000
    x = (rng.random((B, 3, 6, 7)) < 0.3).astype(np.float32)
    x[:, 1] *= (1 - x[:, 0])
    x[:, 2] = rng.integers(0, 2, B).astype(np.float32)[:, None, None]
    eng = B200Net(net, max_batch=4096)
    flags = eng.path_flags()
    assert bool(flags & 4) == ("AZ_PAIR_MIN_ROWS" in env) and bool(flags & 2) == ("AZ_NO_FRONT2" not in env)
    l, v, p = eng.forward_numpy(x)
    with torch.no_grad():
        rl, rv = net(torch.from_numpy(x[:600]))
    check_close(l[:600], v[:600], p[:600], rl.numpy(), rv.numpy())
    eng.close()


@pytest.mark.parametrize("g,n,iters,vl,dim,depth", [(orc.CONNECT4, 256, 200, 1, 128, 1), (orc.GOMOKU, 128, 160, 1, 64, 2),
                                                    (orc.GOMOKU, 96, 200, 4, 64, 1), (orc.TICTACTOE, 64, 100, 1, 64, 1)])
def test_search_with_net_replays_exactly_in_oracle(g, n, iters, vl, dim, depth):
    import alpha_zero_b200 as az
    from alpha_zero_b200.ai.nn import B200Net
    net = make_net(g.rows, g.cols, g.planes, g.action_dim, dim, depth, 3)
    eng = B200Net(net, max_batch=n * vl)
    rng = np.random.default_rng(77)
    cells, players, mcs = random_positions(g, n, rng, 6)
    spec = az.game_spec(g.name, g.rows)
    f = az.Forest(spec, n, iters, vl=vl)
    f.set_roots_words(words_from_cells(g, cells), players, mcs)
    # step by step, logging every evaluation the GPU net produced
    f.begin_move(iters)
    keys, pri, val = [], [], []
    for _ in range(2 * iters + 8):
        f.advance()
        eng.eval_leaves(f)
        r = f.read_leaves()
        if r["n"]:
            k = np.zeros((r["n"], 2 * spec.nw + 1), np.uint64)
            k[:, :2 * spec.nw] = r["words"].reshape(r["n"], -1)
            k[:, -1] = r["players"] if g.planes == 3 else 0
            keys.append(k)
            pri.append(r["priors"])
            val.append(r["values"])
        if f.counts()[1] >= n:
            break
    assert f.counts()[1] == n
    table = co.Table(np.concatenate(keys), np.concatenate(pri), np.concatenate(val))
    ref = co.search_batch(g, cells, players, mcs, iters, vl, table=table)
    assert int(ref["stats"][7]) == 0                              # no table misses: same leaves visited
    compare_forest_with_oracle(f, ref)
    vis_step, q_step = f.root_visits()
    # the CUDA-graph path must give the same search
    f.search_net(eng, iters)
    vis_graph, q_graph = f.root_visits()
    assert np.array_equal(vis_step, vis_graph) and np.array_equal(q_step, q_graph)
    f.close()
    eng.close()


@pytest.mark.parametrize("rows,dim,depth,patch", [(7, 256, 2, 5), (15, 128, 2, 3)])
def test_full_attention_kernels_agree(rows, dim, depth, patch, monkeypatch):
    """The tcgen05 all-token attention (az_attn.cuh) against the scalar kernel it replaces (AZ_NO_ATTN_MMA=1), same net,
    same boards: both must sit inside the torch tolerance and next to each other (P is rounded to bf16 before the P V MMA,
    the scalar kernel keeps it in fp32)."""
    import torch
    from alpha_zero_b200.ai.nn import B200Net
    A = rows * rows
    net = make_net(rows, rows, 2, A, dim, depth, patch)
    rng = np.random.default_rng(4)
    B = 333
    x = (rng.random((B, 2, rows, rows)) < 0.3).astype(np.float32)
    x[:, 1] *= (1 - x[:, 0])
    eng = B200Net(net, max_batch=512)
    assert eng.path_flags() & 32
    l, v, p = eng.forward_numpy(x)
    eng.close()
    monkeypatch.setenv("AZ_NO_ATTN_MMA", "1")
    eng = B200Net(net, max_batch=512)
    assert not eng.path_flags() & 32
    l0, v0, p0 = eng.forward_numpy(x)
    eng.close()
    with torch.no_grad():
        rl, rv = net(torch.from_numpy(x))
    check_close(l, v, p, rl.numpy(), rv.numpy())
    check_close(l0, v0, p0, rl.numpy(), rv.numpy())
    err_new, err_old = np.abs(l - rl.numpy()).max(), np.abs(l0 - rl.numpy()).max()
    print(f"\nlogit error vs torch fp32: tcgen05 attention {err_new:.4f}, scalar attention {err_old:.4f}; between them {np.abs(l - l0).max():.4f}")
    assert np.abs(l - l0).max() <= 0.05


@pytest.mark.parametrize("rows,cols,C,A,dim,patch", [
    (6, 7, 3, 7, 512, 5),          # Connect4, BASELINE configs[1]: cluster of 4
    (7, 7, 2, 49, 512, 5),         # Gomoku 7x7: 49 policy outputs + value in the 64-wide head tile
    (6, 7, 3, 7, 256, 5),          # cluster of 2
    (6, 7, 3, 7, 128, 5),          # one CTA per tile
])
def test_fused_tail_matches_gemm_chain(rows, cols, C, A, dim, patch, monkeypatch):
    """az_tail_kernel (out-proj .. heads of the cls rows in one cluster kernel; opt-in with AZ_FUSED_TAIL=1)
    against the four-GEMM chain and against torch fp32, at batch sizes around the 64-row tile and at more tiles than clusters;
    and a row's answer must not depend on its position in the batch."""
    import torch
    from alpha_zero_b200.ai.nn import B200Net
    net = make_net(rows, cols, C, A, dim, 1, patch)
    rng = np.random.default_rng(9)
    B = 4096
    x = (rng.random((B, C, rows, cols)) < 0.3).astype(np.float32)
    x[:, 1] *= (1 - x[:, 0])
    if C == 3:
        x[:, 2] = rng.integers(0, 2, B).astype(np.float32)[:, None, None]
    monkeypatch.setenv("AZ_FUSED_TAIL", "1")
    eng = B200Net(net, max_batch=4096)
    if not eng.path_flags() & 2:
        eng.close()
        pytest.skip("this shape does not run the reassociated front kernel the fused tail follows")
    assert eng.path_flags() & 64, "the fused tail must be the path that runs"
    l, v, p = eng.forward_numpy(x)
    for nb in (1, 63, 64, 65, 129, 1111):
        ls, vs, ps = eng.forward_numpy(x[:nb])
        assert np.array_equal(ls, l[:nb]) and np.array_equal(vs, v[:nb]) and np.array_equal(ps, p[:nb]), f"batch of {nb}"
    l2, v2, _ = eng.forward_numpy(x[-77:])
    assert np.array_equal(l2, l[-77:]) and np.array_equal(v2, v[-77:])
    eng.close()
    monkeypatch.delenv("AZ_FUSED_TAIL")
    eng = B200Net(net, max_batch=4096)
    assert not eng.path_flags() & 64
    l0, v0, p0 = eng.forward_numpy(x)
    eng.close()
    with torch.no_grad():
        rl, rv = net(torch.from_numpy(x[:600]))
    check_close(l[:600], v[:600], p[:600], rl.numpy(), rv.numpy())
    d = np.abs(l - l0).max()
    print(f"\nfused tail vs GEMM chain: max |dlogit| {d:.5f}, max |dvalue| {np.abs(v - v0).max():.5f}; vs torch fp32: fused {np.abs(l[:600] - rl.numpy()).max():.4f}, chain {np.abs(l0[:600] - rl.numpy()).max():.4f}")
    assert d <= 0.03 and np.abs(p - p0).max() <= 0.01


@pytest.mark.parametrize("rows,cols,C,A,dim,patch", [(6, 7, 3, 7, 512, 5), (7, 7, 2, 49, 512, 5), (6, 7, 3, 7, 256, 5)])
def test_chain_kernel_is_bit_identical_to_separate_launches(rows, cols, C, A, dim, patch, monkeypatch):
    """az_chain_kernel (csrc/az_chain.cuh) runs the four cls-row GEMMs of small evaluation batches back to back in one
    persistent kernel with grid-wide barriers between the layers: same tiles, same epilogues, so logits, values and priors
    must equal those of the four separate launches bit for bit, at every batch size and on repeated calls (the barrier
    counters are reset by the last CTA of each launch).  Opt-in with AZ_CHAIN=1: slower inside the search graph."""
    from alpha_zero_b200.ai.nn import B200Net
    net = make_net(rows, cols, C, A, dim, 1, patch)
    rng = np.random.default_rng(11)
    B = 4096
    x = (rng.random((B, C, rows, cols)) < 0.3).astype(np.float32)
    x[:, 1] *= (1 - x[:, 0])
    if C == 3:
        x[:, 2] = rng.integers(0, 2, B).astype(np.float32)[:, None, None]
    monkeypatch.setenv("AZ_CHAIN", "1")
    eng = B200Net(net, max_batch=4096)
    if not eng.path_flags() & 2:
        eng.close()
        pytest.skip("this shape does not run the reassociated front kernel")
    if dim == 512:
        assert eng.path_flags() & 256, "the chain kernel must be the path that runs"
    elif not eng.path_flags() & 256:
        eng.close()
        pytest.skip("tiles of 64 columns were chosen for this width: separate launches")
    outs = {}
    for nb in (1, 127, 128, 129, 1111, 1792, 4096, 1792, 5):
        outs[nb] = eng.forward_numpy(x[:nb])
    eng.close()
    monkeypatch.delenv("AZ_CHAIN")
    eng = B200Net(net, max_batch=4096)
    assert not eng.path_flags() & 256
    for nb, (l, v, p) in outs.items():
        l0, v0, p0 = eng.forward_numpy(x[:nb])
        assert np.array_equal(l, l0) and np.array_equal(v, v0) and np.array_equal(p, p0), f"batch of {nb}"
    eng.close()
