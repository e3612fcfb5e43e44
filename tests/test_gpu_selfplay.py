"""Device-side self-play (csrc/az_selfplay.cu) against the oracle.

The engine logs, per recorded position, the uniform that drew the move; the Dirichlet noise each search used is read
back before the search.  Every game is then replayed in the C oracle: same noise -> the visit counts and root Q must be
bit-exact, the move must be what np.random.choice's inverse CDF (searchsorted side='right') returns for the logged
uniform -- or the first most-visited child in CHILD order (node.py:76-81) on argmax plies --, and plies / winner must
follow from the oracle's rules.  Also: trajectories do not depend on slot count or on how games are sharded over
engines (the multi-GPU invariant), the device Dirichlet sampler has the right distribution, NCCL gather on all GPUs.
"""
import os
import socket
import sys

import numpy as np
import pytest

import az_oracle as orc
import c_oracle as co

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def game_cls(g):
    from alpha_zero_b200 import games
    G = {"tictactoe": games.TicTacToe, "connect4": games.Connect4, "gomoku": games.Gomoku}[g.name]
    if g.name == "gomoku":
        G.rows = G.cols = g.rows
        G.action_dim = G.state_dim = g.rows * g.cols
    return G


def _step_by_step(g, n_games, slots, iters, vl, seed, hseed, qbits, sample_plies, noise):
    """Drive the engine ply by ply, logging the noise each (game, ply) searched with."""
    import alpha_zero_b200 as az
    f = az.Forest(az.game_spec(g.name, g.rows), slots, iters, vl=vl)
    sp = az.SelfPlayEngine(f, n_games, seed, sample_plies, noise=noise)
    sp.begin()
    ply_of = np.zeros(n_games, np.int64)
    noise_log = {}
    for _ in range(n_games * g.rows * g.cols + 4):
        sp.prepare()
        sg, nz = sp.slots(with_noise=True)
        for s, k in enumerate(sg):
            if k >= 0:
                noise_log[(int(k), int(ply_of[k]))] = nz[s].copy()
                ply_of[k] += 1
        f.search_hash(iters, hseed, qbits)
        sp.move()
        if sp.status()["live"] == 0:
            break
    st = sp.status()
    assert st["live"] == 0 and st["finished"] == n_games and st["started"] == n_games
    arr = sp.trajectory_arrays(with_uniforms=True)
    assert st["positions"] == int(arr["plies"].sum())
    sp.close()
    f.close()
    return arr, noise_log


def _replay_in_oracle(g, arr, noise_log, iters, vl, hseed, qbits, sample_plies, noise):
    n_games = len(arr["plies"])
    ncell = g.rows * g.cols
    cells = np.zeros((n_games, ncell), np.uint8)
    alive = np.ones(n_games, bool)
    winner = np.full(n_games, -2)
    for ply in range(ncell):
        idx = np.nonzero(alive)[0]
        if not len(idx):
            break
        nz = np.stack([noise_log[(int(k), ply)] for k in idx]) if noise else None
        ref = co.search_batch(g, cells[idx], np.full(len(idx), ply & 1, np.int32), np.full(len(idx), ply, np.int32), iters, vl,
                              seed=hseed, qbits=qbits, noise=nz)
        for j, k in enumerate(idx):
            nc = int(ref["n_children"][j])
            ch_cells, ch_vis = ref["cells"][j, :nc], ref["visits"][j, :nc]
            vis = np.zeros(g.action_dim, np.int64)
            ai = ch_cells % g.cols if g.gid == 1 else ch_cells
            vis[ai] = ch_vis
            # board before the move, visit counts, root Q: bit-exact
            w = arr["words"][k, ply]
            got = np.zeros(ncell, np.uint8)
            for side in (0, 1):
                for wi in range(w.shape[1]):
                    for bit in range(64):
                        c = wi * 64 + bit
                        if c < ncell and (int(w[side, wi]) >> bit) & 1:
                            got[c] = side + 1
            assert np.array_equal(got, cells[k]), f"game {k} ply {ply}: recorded board"
            assert np.array_equal(arr["visits"][k, ply].astype(np.int64), vis), f"game {k} ply {ply}: visits"
            assert arr["q"][k, ply] == ref["root_value"][j] / ref["root_visit"][j], f"game {k} ply {ply}: root Q"
            u = arr["u"][k, ply]
            if ply < sample_plies:
                assert 0.0 <= u < 1.0
                pi = vis / np.sum(vis.astype(np.float64))                # utils.py:46-55
                cdf = pi.cumsum()
                cdf /= cdf[-1]
                a = int(cdf.searchsorted(u, side="right"))               # np.random.choice (node.py:92)
                cell = int(ch_cells[list(ai).index(a)])
            else:
                assert u == -1.0
                cell = int(ch_cells[int(np.argmax(ch_vis))])             # first maximum in child order (node.py:76-81)
            assert int(arr["action"][k, ply]) == cell, f"game {k} ply {ply}: move"
            cells[k, cell] = (ply & 1) + 1
            wv = co.winner(g, cells[k], ply & 1, cell)
            if wv != -1 or ply + 1 == ncell:
                alive[k] = False
                winner[k] = wv
                assert int(arr["plies"][k]) == ply + 1, f"game {k}: plies"
    assert not alive.any()
    assert np.array_equal(winner, arr["winner"])
    assert (arr["first"] == 0).all()


@pytest.mark.parametrize("name,n_games,slots,iters,vl", [("tictactoe", 24, 8, 40, 1), ("connect4", 20, 8, 64, 1), ("gomoku", 16, 6, 48, 1),
                                                        ("gomoku", 12, 12, 64, 4)])
def test_device_selfplay_replays_in_oracle(name, n_games, slots, iters, vl):
    g = {"tictactoe": orc.TICTACTOE, "connect4": orc.CONNECT4, "gomoku": orc.GOMOKU}[name]
    sample_plies = 8 if name == "gomoku" else g.rows * g.cols + 1
    arr, noise_log = _step_by_step(g, n_games, slots, iters, vl, seed=20240 + n_games, hseed=5, qbits=12, sample_plies=sample_plies, noise=True)
    _replay_in_oracle(g, arr, noise_log, iters, vl, 5, 12, sample_plies, True)


def test_argmax_ties_take_first_child_in_child_order():
    """qbits = 2 makes equal visit counts common; no sampling, no noise: every move is max_visit_child."""
    g = orc.GOMOKU
    arr, noise_log = _step_by_step(g, 12, 12, 24, 1, seed=3, hseed=9, qbits=2, sample_plies=0, noise=False)
    _replay_in_oracle(g, arr, noise_log, 24, 1, 9, 2, 0, False)


def test_trajectories_do_not_depend_on_slots_or_sharding():
    """Game g's randomness is keyed by its GLOBAL id: 1 engine x 16 games == 2 engines (ids 0,2,.. / 1,3,..) == fewer slots."""
    from alpha_zero_b200.ai import HashEvaluator
    from alpha_zero_b200.selfplay import play_games_device
    g = orc.CONNECT4
    G = game_cls(g)
    ev = HashEvaluator(11, 10)
    whole = play_games_device(G, ev, 16, 40, seed=77, return_arrays=True)
    few = play_games_device(G, ev, 16, 40, seed=77, concurrent=5, return_arrays=True)
    even = play_games_device(G, ev, 8, 40, seed=77, id_first=0, id_stride=2, return_arrays=True)
    odd = play_games_device(G, ev, 8, 40, seed=77, id_first=1, id_stride=2, return_arrays=True)
    for k in ("plies", "winner", "words", "visits", "q", "action"):
        assert np.array_equal(whole[k], few[k]), k
        assert np.array_equal(whole[k][0::2], even[k]), k
        assert np.array_equal(whole[k][1::2], odd[k]), k
    assert whole["plies"].min() >= 7


def test_play_games_returns_reference_tuples_and_fills_buffer():
    """train.play_games / collect_data on the device engine: tuples in the reference's layout, replayable under the host rules."""
    import torch
    from alpha_zero_b200.ai import Net
    from alpha_zero_b200.replay_buffer import ReplayBuffer
    from alpha_zero_b200 import train
    torch.manual_seed(2)
    np.random.seed(2)
    for g in (orc.GOMOKU, orc.CONNECT4):
        G = game_cls(g)
        net = Net((G.rows, G.cols), 3, 64, G.action_dim, 8, 1, G.feature_dim, dropout=0.1)
        iters = 24
        games, visits = train.play_games(G, net, 20, iters, dirichlet=True, return_visits=True, concurrent=8)
        assert len(games) == 20
        for (boards, actions, policies, qs, winner), vis in zip(games, visits):
            assert len(boards) == len(policies) == len(qs) == len(actions) - 1 == len(vis) >= 1
            b = G().board.copy()
            player, w = 0, -1
            for i, mv in enumerate(actions[1:]):
                assert boards[i].dtype == np.float32 and np.array_equal(boards[i], b)
                assert tuple(mv) in [tuple(m) for m in G.get_valid_moves(b)]
                assert vis[i].sum() == iters - 1 and np.array_equal(policies[i], vis[i] / vis[i].sum())
                assert vis[i][G.get_action_idx(mv)] > 0
                mover = player
                player = G.make_move(b, mover, mv)
                w = G.check_winner(b, mover, mv)
                assert (w != -1 or i + 1 == G.state_dim) == (i == len(boards) - 1)
            assert winner == w
        buf = ReplayBuffer(100000)
        res = train.collect_data(G, net, buf, iterations=8, mcts_iter=16)
        assert sum(res) == 8 and buf.size() >= 8 * 5


def test_device_dirichlet_noise_distribution():
    """np.random.dirichlet([0.03] * A) drawn on the device (Marsaglia-Tsang + boost): rows sum to 1, component mean 1/A,
    variance alpha(alpha0 - alpha) / (alpha0^2 (alpha0 + 1)), and the share of dominant components matches NumPy's sampler."""
    import alpha_zero_b200 as az
    g = orc.GOMOKU
    n = 4096
    f = az.Forest(az.game_spec(g.name, g.rows), n, 8, vl=1)
    sp = az.SelfPlayEngine(f, n, 99, 8, noise=True)
    sp.begin()
    sp.prepare()
    _sg, nz = sp.slots(with_noise=True)
    sp.close()
    f.close()
    A, alpha = g.action_dim, 0.03
    assert np.all(nz >= 0) and np.allclose(nz.sum(axis=1), 1.0, atol=1e-12)
    a0 = A * alpha
    var = alpha * (a0 - alpha) / (a0 * a0 * (a0 + 1))
    assert abs(nz.mean() - 1.0 / A) < 1e-12
    assert abs(nz.var() / var - 1.0) < 0.05
    ref = np.random.default_rng(0).dirichlet([alpha] * A, size=n)
    for thr in (0.5, 0.9, 0.01):
        assert abs((nz > thr).mean() - (ref > thr).mean()) < 0.15 * (ref > thr).mean() + 1e-4
    assert len(np.unique(nz[:, 0])) > n // 2                    # streams differ per game


# ---- NCCL: every visible GPU plays its shard, the gathered buffers are identical and equal to the 1-rank result ------
def _nccl_worker(rank, world, port, out):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import hashlib
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    from alpha_zero_b200 import parallel
    from alpha_zero_b200.ai import HashEvaluator
    from alpha_zero_b200.games import Gomoku
    from alpha_zero_b200.replay_buffer import ReplayBuffer
    buf = ReplayBuffer(10 ** 6)
    res = parallel.collect_data_distributed(Gomoku, HashEvaluator(4, 12), buf, 21, 32, seed=1234, concurrent=4)
    h = hashlib.sha256()
    for s, p, z in buf.buffer:
        h.update(np.ascontiguousarray(s, np.float32).tobytes())
        h.update(np.ascontiguousarray(p, np.float64).tobytes())
        h.update(np.asarray(z, np.float64).tobytes())
    out[rank] = (h.hexdigest(), buf.size(), tuple(res))
    dist.barrier()
    dist.destroy_process_group()


def test_nccl_gather_fills_identical_buffers_on_all_gpus():
    import torch
    import torch.multiprocessing as mp
    world = torch.cuda.device_count()
    if world < 2:
        pytest.skip("needs >= 2 GPUs")
    game_cls(orc.GOMOKU)
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    out = mp.Manager().dict()
    mp.spawn(_nccl_worker, args=(world, port, out), nprocs=world, join=True)
    assert len(set(out.values())) == 1, dict(out)
    # single-process result with the same seed
    import hashlib
    from alpha_zero_b200 import parallel
    from alpha_zero_b200.ai import HashEvaluator
    from alpha_zero_b200.games import Gomoku
    from alpha_zero_b200.replay_buffer import ReplayBuffer
    buf = ReplayBuffer(10 ** 6)
    res = parallel.collect_data_distributed(Gomoku, HashEvaluator(4, 12), buf, 21, 32, seed=1234, concurrent=4)
    h = hashlib.sha256()
    for s_, p, z in buf.buffer:
        h.update(np.ascontiguousarray(s_, np.float32).tobytes())
        h.update(np.ascontiguousarray(p, np.float64).tobytes())
        h.update(np.asarray(z, np.float64).tobytes())
    assert (h.hexdigest(), buf.size(), tuple(res)) == out[0]


# ---- NCCL, config 5: device buffer filled from the gathered games, gradient all-reduce keeps the replicas identical ----
def _device_buffer_digest(dbuf):
    import hashlib
    h = hashlib.sha256()
    for x in dbuf.read():
        h.update(np.ascontiguousarray(x.cpu().numpy()).tobytes())
    return h.hexdigest()


def _nccl_train_worker(rank, world, port, out):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    from alpha_zero_b200 import parallel, train
    from alpha_zero_b200.ai import HashEvaluator, Net
    from alpha_zero_b200.games import Gomoku
    from alpha_zero_b200.replay_buffer import DeviceReplayBuffer
    dbuf = DeviceReplayBuffer(Gomoku, 100000, seed=50 + rank)         # same content everywhere, different batches per rank
    res = parallel.collect_data_device(Gomoku, HashEvaluator(4, 12), dbuf, 21, 32, seed=1234, concurrent=4)
    digest = _device_buffer_digest(dbuf)
    torch.manual_seed(0)
    net = Net((7, 7), 3, 64, 49, 8, 1, 2, dropout=0.0)
    before = torch.cat([p.detach().reshape(-1) for p in net.parameters()]).clone()
    train.train(net, 32, dbuf, 3, 1e-3, "cuda")
    flat = torch.cat([p.detach().reshape(-1) for p in net.parameters()])
    gathered = [torch.empty_like(flat) for _ in range(world)]
    dist.all_gather(gathered, flat)
    same = all(torch.equal(gathered[0], t) for t in gathered)
    moved = not torch.equal(flat.cpu(), before)
    out[rank] = (digest, dbuf.size(), tuple(res), same, moved, train.last_stats.get("buckets", 0) >= 1)
    dist.barrier()
    dist.destroy_process_group()


def test_nccl_device_buffer_and_gradient_all_reduce():
    import torch
    import torch.multiprocessing as mp
    world = torch.cuda.device_count()
    if world < 2:
        pytest.skip("needs >= 2 GPUs")
    game_cls(orc.GOMOKU)
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    out = mp.Manager().dict()
    mp.spawn(_nccl_train_worker, args=(world, port, out), nprocs=world, join=True)
    assert len({v[:3] for v in out.values()}) == 1, dict(out)       # identical buffers and results on every rank
    assert all(v[3] and v[4] and v[5] for v in out.values()), dict(out)   # replicas identical, weights moved, buckets reduced
    from alpha_zero_b200 import parallel
    from alpha_zero_b200.ai import HashEvaluator
    from alpha_zero_b200.games import Gomoku
    from alpha_zero_b200.replay_buffer import DeviceReplayBuffer
    dbuf = DeviceReplayBuffer(Gomoku, 100000, seed=1)               # one process, same seed: the same buffer (rank-count independent)
    res = parallel.collect_data_device(Gomoku, HashEvaluator(4, 12), dbuf, 21, 32, seed=1234, concurrent=4)
    assert (_device_buffer_digest(dbuf), dbuf.size(), tuple(res)) == out[0][:3]
    dbuf.close()


def test_collect_data_device_single_gpu_matches_host_buffer():
    """Same games through the host path (collect_data_distributed -> save_data_to_buffer) and the device path."""
    from alpha_zero_b200 import parallel
    from alpha_zero_b200.ai import HashEvaluator
    from alpha_zero_b200.replay_buffer import DeviceReplayBuffer, ReplayBuffer
    for g in (orc.GOMOKU, orc.CONNECT4):
        G = game_cls(g)
        dbuf = DeviceReplayBuffer(G, 100000, seed=1)
        r1 = parallel.collect_data_device(G, HashEvaluator(7, 10), dbuf, 12, 24, seed=99, concurrent=5)
        hbuf = ReplayBuffer(100000)
        r2 = parallel.collect_data_distributed(G, HashEvaluator(7, 10), hbuf, 12, 24, seed=99, concurrent=5)
        assert list(r1) == list(r2) and dbuf.size() == hbuf.size()
        s, p, z = (x.cpu().numpy() for x in dbuf.read())
        assert np.array_equal(s, np.stack([e[0] for e in hbuf.buffer]))
        assert np.array_equal(p, np.stack([e[1] for e in hbuf.buffer]).astype(np.float32))
        assert np.array_equal(z[:, 0], np.asarray([e[2][0] for e in hbuf.buffer], np.float32))
        dbuf.close()


@pytest.mark.parametrize("name,vl,rounds", [("connect4", 1, 1), ("connect4", 1, 4), ("gomoku", 1, 3), ("gomoku", 8, 1), ("tictactoe", 1, 2)])
def test_continuous_plies_give_the_lockstep_trajectories(name, vl, rounds):
    """az_selfplay_run_async_*: slots move as soon as their own search is finished.  A game's trajectory is a function of
    (game id, ply) only, so it must equal the lockstep run bit for bit -- with several cached simulations per launch too."""
    from alpha_zero_b200.ai import HashEvaluator, MCTS
    from alpha_zero_b200.selfplay import play_games_device
    g = {"tictactoe": orc.TICTACTOE, "connect4": orc.CONNECT4, "gomoku": orc.GOMOKU}[name]
    G = game_cls(g)
    ev = HashEvaluator(13, 10)
    iters = 48
    lock = play_games_device(G, ev, 30, iters, vl=vl, seed=5, concurrent=7, return_arrays=True, lockstep=True)
    cont = play_games_device(G, ev, 30, iters, vl=vl, seed=5, concurrent=7, return_arrays=True, rounds=rounds)
    for k in ("plies", "winner", "first", "words", "visits", "q", "action"):
        assert np.array_equal(lock[k], cont[k]), k
    assert (lock["plies"] > 0).all()


def test_continuous_plies_with_the_net_match_lockstep():
    import torch
    from alpha_zero_b200.ai import Net, MCTS
    from alpha_zero_b200.selfplay import play_games_device
    G = game_cls(orc.CONNECT4)
    torch.manual_seed(3)
    net = Net((G.rows, G.cols), 5, 128, G.action_dim, 8, 1, G.feature_dim, dropout=0.0).eval()
    lock = play_games_device(G, net, 40, 64, seed=8, concurrent=16, return_arrays=True, lockstep=True)
    cont = play_games_device(G, net, 40, 64, seed=8, concurrent=16, return_arrays=True)          # default: 4 cached simulations per launch
    for k in ("plies", "winner", "words", "visits", "q", "action"):
        assert np.array_equal(lock[k], cont[k]), k
