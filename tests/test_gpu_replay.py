"""Device-resident replay buffer and the fused training loss (csrc/az_replay.cu) against the reference's data path:
train.save_data_to_buffer / rotate_data / flip_data (train.py:8-49), ReplayBuffer.sample (replay_buffer.py:15-25) and
the loss of train.train (train.py:101-112)."""
import numpy as np
import pytest

import az_oracle as orc
from golden_util import load_json, spec_of, GOLDEN

pytestmark = pytest.mark.gpu


def game_cls(g):
    from alpha_zero_b200 import games
    G = {"tictactoe": games.TicTacToe, "connect4": games.Connect4, "gomoku": games.Gomoku}[g.name]
    if g.name == "gomoku":
        G.rows = G.cols = g.rows
        G.action_dim = G.state_dim = g.rows * g.cols
    return G


def _to_cuda(arr):
    import torch
    return dict(plies=torch.from_numpy(arr["plies"]).cuda(), winner=torch.from_numpy(arr["winner"]).cuda(),
                first=torch.from_numpy(arr["first"]).cuda(), words=torch.from_numpy(arr["words"].view(np.int64)).cuda(),
                visits=torch.from_numpy(arr["visits"].view(np.int16)).cuda())


def _host_buffer(G, spec, arr, size):
    from alpha_zero_b200.replay_buffer import ReplayBuffer
    from alpha_zero_b200.selfplay import games_from_arrays
    from alpha_zero_b200 import parallel
    buf = ReplayBuffer(size)
    parallel.store_games(G, buf, games_from_arrays(spec, arr))
    return buf


def _same_entries(dev, host):
    s, p, z = (x.cpu().numpy() for x in dev.read())
    assert len(s) == host.size()
    hs = np.stack([e[0] for e in host.buffer])
    hp = np.stack([e[1] for e in host.buffer]).astype(np.float32)      # the cast ReplayBuffer.sample applies
    hz = np.asarray([e[2][0] for e in host.buffer], np.float32)
    assert np.array_equal(s, hs) and np.array_equal(p, hp) and np.array_equal(z[:, 0], hz)


def test_reference_buffer_goldens():
    """Trajectories recorded from the UNMODIFIED reference -> device buffer == the reference's own buffer contents."""
    import alpha_zero_b200 as az
    from alpha_zero_b200.forest import planes_to_words
    from alpha_zero_b200.replay_buffer import DeviceReplayBuffer
    recs = load_json("selfplay.json")
    arrs = np.load(f"{GOLDEN}/selfplay.npz")
    checked = 0
    for i, rec in enumerate(recs):
        if f"t{i}_buf_states" not in arrs:
            continue
        g = spec_of(rec)
        G = game_cls(g)
        spec = az.game_spec(g.name, g.rows)
        boards, pol = arrs[f"t{i}_boards"], arrs[f"t{i}_policies"]
        plies, P, A = len(boards), spec.state_dim, spec.action_dim
        arr = dict(plies=np.array([plies], np.int32), winner=np.array([rec["winner"]], np.int32), first=np.zeros(1, np.int32),
                   words=np.zeros((1, P, 2, spec.nw), np.uint64), visits=np.zeros((1, P, A), np.uint16))
        arr["words"][0, :plies] = planes_to_words(spec, boards)
        vis = np.rint(pol * (rec["iters"] - 1)).astype(np.int64)
        assert np.array_equal(vis / vis.sum(axis=1, keepdims=True), pol)     # the golden policies are visit counts / sum
        arr["visits"][0, :plies] = vis
        dev = DeviceReplayBuffer(G, 100000, seed=1)
        n = dev.add_trajectories(_to_cuda(arr))
        s, p, z = (x.cpu().numpy() for x in dev.read())
        assert n == len(arrs[f"t{i}_buf_states"])
        assert np.array_equal(s, arrs[f"t{i}_buf_states"])
        assert np.array_equal(p, arrs[f"t{i}_buf_policies"].astype(np.float32))
        assert np.array_equal(z[:, 0], arrs[f"t{i}_buf_z"].reshape(-1).astype(np.float32))
        dev.close()
        checked += 1
    assert checked >= 2
    game_cls(orc.GOMOKU)


@pytest.mark.parametrize("name", ["gomoku", "connect4", "tictactoe"])
def test_device_buffer_equals_host_path_and_wraps_like_a_deque(name):
    import alpha_zero_b200 as az
    from alpha_zero_b200.ai import HashEvaluator
    from alpha_zero_b200.replay_buffer import DeviceReplayBuffer
    from alpha_zero_b200.selfplay import play_games_device
    g = {"tictactoe": orc.TICTACTOE, "connect4": orc.CONNECT4, "gomoku": orc.GOMOKU}[name]
    G = game_cls(g)
    spec = az.game_spec(g.name, g.rows)
    arr = play_games_device(G, HashEvaluator(3, 10), 24, 32, seed=5, concurrent=10, return_arrays=True)
    for size in (10 ** 6, 301):                                      # roomy; smaller than one batch of games (oldest dropped)
        dev = DeviceReplayBuffer(G, size, seed=2)
        dev.add_trajectories(_to_cuda(arr))
        _same_entries(dev, _host_buffer(G, spec, arr, size))
        dev.close()
    dev = DeviceReplayBuffer(G, 500, seed=2)                         # several adds: ring wraps across calls
    host = _host_buffer(G, spec, arr, 500)
    dev.add_trajectories(_to_cuda(arr))
    half = {k: v[:11] for k, v in arr.items()}
    dev.add_trajectories(_to_cuda(half))
    from alpha_zero_b200 import parallel
    from alpha_zero_b200.selfplay import games_from_arrays
    parallel.store_games(G, host, games_from_arrays(spec, half))
    _same_entries(dev, host)
    # sampling: distinct entries, each equal to the entry it names
    B = 256
    s, p, z, idx = dev.sample(B, return_indices=True)
    idx = idx.cpu().numpy()
    assert len(np.unique(idx)) == B and idx.min() >= 0 and idx.max() < dev.size()
    rs, rp, rz = (x.cpu().numpy() for x in dev.read())
    assert np.array_equal(s.cpu().numpy(), rs[idx]) and np.array_equal(p.cpu().numpy(), rp[idx]) and np.array_equal(z.cpu().numpy(), rz[idx])
    assert tuple(s.shape) == (B, spec.planes, spec.rows, spec.cols) and tuple(p.shape) == (B, spec.action_dim) and tuple(z.shape) == (B, 1)
    s2, _p2, _z2, idx2 = dev.sample(B, return_indices=True)
    assert not np.array_equal(idx, idx2.cpu().numpy())               # a new draw every call
    counts = np.zeros(dev.size())
    for _ in range(200):
        counts[dev.sample(50, return_indices=True)[3].cpu().numpy()] += 1
    expect = 200 * 50 / dev.size()
    assert counts.min() > 0 and abs(counts.mean() - expect) < 1e-9 and counts.std() < 3.0 * np.sqrt(expect)
    with pytest.raises(az.lib.AzError):
        dev.sample(dev.size() + 1)
    hb = dev.to_host()
    assert hb.size() == dev.size()
    dev.close()


def test_fused_loss_matches_reference_formula():
    """train.py:110-112: policy_loss = mean(sum(-pi * log_softmax(logits))), value_loss = mse; values and gradients to 1e-5."""
    import torch
    from alpha_zero_b200.train import policy_value_loss
    torch.manual_seed(0)
    for B, A in ((512, 49), (33, 7), (64, 225)):
        logits = (torch.randn(B, A, device="cuda") * 3).requires_grad_()
        values = torch.tanh(torch.randn(B, 1, device="cuda")).requires_grad_()
        pis = torch.softmax(torch.randn(B, A, device="cuda") * 2, dim=1)
        pis[:, 0] = 0
        pis = pis / pis.sum(dim=1, keepdim=True)
        zs = torch.randint(-1, 2, (B, 1), device="cuda").float()
        pl, vl = policy_value_loss(logits, values, pis, zs)
        (pl + 0.5 * vl).backward()
        l2, v2 = logits.detach().clone().requires_grad_(), values.detach().clone().requires_grad_()
        rpl = torch.mean(torch.sum(-pis * torch.log_softmax(l2, dim=1), dim=1))
        rvl = torch.nn.functional.mse_loss(v2, zs)
        (rpl + 0.5 * rvl).backward()
        assert abs(pl.item() - rpl.item()) <= 1e-5 * max(1.0, abs(rpl.item())) and abs(vl.item() - rvl.item()) <= 1e-5
        assert torch.allclose(logits.grad, l2.grad, rtol=1e-5, atol=1e-7) and torch.allclose(values.grad, v2.grad, rtol=1e-5, atol=1e-8)


def test_train_step_on_device_buffer_matches_reference_step():
    """One Adam step of train.train fed by the device buffer == the reference's step (train.py:93-117) on the same batch."""
    import copy
    import torch
    from alpha_zero_b200 import train
    from alpha_zero_b200.ai import Net, HashEvaluator
    from alpha_zero_b200.replay_buffer import DeviceReplayBuffer
    from alpha_zero_b200.selfplay import play_games_device
    G = game_cls(orc.GOMOKU)
    arr = play_games_device(G, HashEvaluator(8, 10), 16, 32, seed=9, return_arrays=True)
    dev = DeviceReplayBuffer(G, 10 ** 5, seed=4)
    dev.add_trajectories(_to_cuda(arr))
    torch.manual_seed(1)
    net = Net((G.rows, G.cols), 3, 64, G.action_dim, 8, 1, G.feature_dim, dropout=0.0)
    ref = copy.deepcopy(net).to("cuda").train()
    dev.draws = 0
    out = train.train(net, 64, dev, 1, 2.5e-4, "cuda")
    dev.draws = 0
    states, pis, zs = dev.sample(64)                                 # the same draw again
    opt = torch.optim.Adam(ref.parameters(), lr=2.5e-4)
    l2 = sum(torch.sum(p ** 2) for name, p in ref.named_parameters() if "bias" not in name and "LayerNorm" not in name)
    logits, values = ref(states)
    pl = torch.mean(torch.sum(-pis * torch.log_softmax(logits, dim=1), dim=1))
    vl = torch.nn.functional.mse_loss(values, zs)
    loss = pl + vl + 1e-4 * l2
    opt.zero_grad()
    loss.backward()
    opt.step()
    assert abs(out[0] - loss.item()) <= 1e-5 * abs(loss.item()) and abs(out[1] - pl.item()) <= 1e-5 and abs(out[2] - vl.item()) <= 1e-5
    for (name, a), b in zip(net.named_parameters(), ref.parameters()):
        # gradients to 1e-5 relative (atol: gradients that are zero in theory, e.g. the key bias, are rounding noise on both sides)
        assert torch.allclose(a.grad, b.grad, rtol=1e-5, atol=2e-7), name
        assert (a - b).abs().max().item() <= 1e-5, name               # Adam's first step is lr * g / (|g| + eps): bounded by lr = 2.5e-4
    dev.close()
