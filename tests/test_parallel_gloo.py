"""The N > 1 path on CPU: world_size-2 `gloo` processes shard games, all-gather the compact
trajectory arrays (the layout of the device trajectory store, csrc/az_selfplay.cu) and must end with
IDENTICAL replay buffers, equal to the single-process buffer built in global game order; the gradient
all-reduce of train.train keeps the replicas in lock step.  (The searches themselves need a GPU; here
every rank fabricates its games with random legal play through the host Game classes, which exercises
gather / unpack / canonicalisation / augmentation end to end.  tests/test_gpu_selfplay.py runs the same
exchange over NCCL on every visible GPU.)"""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _fake_game(Game, gid, iters=50):
    """One random game in the format of train.play_games, deterministic in the global id."""
    rng = np.random.default_rng(1000 + gid)
    g = Game()
    board, player, ply = g.board, 0, 0
    boards, actions, policies, qs, visits = [], [(-1, -1)], [], [], []
    winner = -1
    while True:
        legal = Game.get_valid_moves(board)
        vis = np.zeros(Game.action_dim, np.int64)
        for _ in range(iters - 1):
            vis[Game.get_action_idx(legal[rng.integers(len(legal))])] += 1
        visits.append(vis)
        policies.append(vis / vis.sum())
        boards.append(board.copy())
        qs.append(float(rng.standard_normal()))
        mv = legal[int(rng.integers(len(legal)))]
        nxt = Game.make_move(board, player, mv)
        actions.append(mv)
        ply += 1
        w = Game.check_winner(board, player, mv)
        if w != -1 or ply == Game.state_dim:
            winner = w
            break
        player = nxt
    return (boards, actions, policies, qs, winner), np.stack(visits)


def _arrays_of(Game, fakes):
    """Fabricated games -> the compact arrays of selfplay.SelfPlayEngine.trajectory_arrays."""
    from alpha_zero_b200.forest import game_spec, planes_to_words
    g = game_spec(Game.name, Game.rows)
    n, P, A = len(fakes), g.state_dim, g.action_dim
    arr = dict(plies=np.zeros(n, np.int32), winner=np.zeros(n, np.int32), first=np.zeros(n, np.int32),
               words=np.zeros((n, P, 2, g.nw), np.uint64), visits=np.zeros((n, P, A), np.uint16), q=np.zeros((n, P)),
               action=np.zeros((n, P), np.int16))
    for k, ((boards, actions, _pol, qs, winner), vis) in enumerate(fakes):
        p = len(boards)
        arr["plies"][k], arr["winner"][k] = p, winner
        arr["words"][k, :p] = planes_to_words(g, np.stack(boards))
        arr["visits"][k, :p] = vis
        arr["q"][k, :p] = qs
        arr["action"][k, :p] = [r * g.cols + c for r, c in actions[1:]]
    return g, arr


def _buffer_digest(buffer):
    import hashlib
    h = hashlib.sha256()
    for s, p, z in buffer.buffer:
        h.update(np.ascontiguousarray(s, np.float32).tobytes())
        h.update(np.ascontiguousarray(p, np.float64).tobytes())
        h.update(np.asarray(z, np.float64).tobytes())
    return h.hexdigest(), buffer.size()


def _worker(rank, world, port, game_name, n_games, out):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from alpha_zero_b200 import games, parallel
    from alpha_zero_b200.replay_buffer import ReplayBuffer
    from alpha_zero_b200.ai.nn import Net
    from alpha_zero_b200 import train as tr
    G = {"gomoku": games.Gomoku, "connect4": games.Connect4}[game_name]
    ids = parallel.shard_games(n_games)
    assert ids == list(range(rank, n_games, world))
    from alpha_zero_b200.selfplay import games_from_arrays
    spec, local = _arrays_of(G, [_fake_game(G, g) for g in ids])
    merged = parallel.trajectories_to_host(parallel.gather_trajectories(local, n_games))
    _spec, whole = _arrays_of(G, [_fake_game(G, g) for g in range(n_games)])
    for k in parallel.TRAJ_KEYS:                                     # global game order, whatever the rank count
        assert np.array_equal(merged[k], whole[k]), k
    buf = ReplayBuffer(100000)
    results = parallel.store_games(G, buf, games_from_arrays(spec, merged))
    digest = _buffer_digest(buf)
    # gradient all-reduce: different samples per rank, identical parameters afterwards
    torch.manual_seed(7)
    net = Net((G.rows, G.cols), 3, 64, G.action_dim, 4, 1, G.feature_dim, dropout=0.0)
    np.random.seed(100 + rank)
    tr.train(net, 16, buf, 2, 1e-3, "cpu")
    flat = torch.cat([p.detach().reshape(-1) for p in net.parameters()])
    gathered = [torch.empty_like(flat) for _ in range(world)]
    dist.all_gather(gathered, flat)
    same = all(torch.equal(gathered[0], t) for t in gathered)
    out[rank] = (digest, results, same)
    dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("game_name,n_games", [("gomoku", 7), ("connect4", 5)])
def test_two_ranks_build_identical_buffers(game_name, n_games):
    from alpha_zero_b200 import games, parallel
    from alpha_zero_b200.replay_buffer import ReplayBuffer
    G = {"gomoku": games.Gomoku, "connect4": games.Connect4}[game_name]
    ref = ReplayBuffer(100000)
    single = [_fake_game(G, g) for g in range(n_games)]
    ref_results = [0, 0, 0]
    from alpha_zero_b200.train import save_data_to_buffer
    for (boards, actions, policies, qs, winner), _ in single:
        ref_results[2 if winner == -1 else winner] += 1
        save_data_to_buffer(G, ref, (boards, actions, policies, qs, winner, 0 if winner == -1 else 1))
    mgr = mp.Manager()
    out = mgr.dict()
    port = _free_port()
    mp.spawn(_worker, args=(2, port, game_name, n_games, out), nprocs=2, join=True)
    assert out[0][0] == out[1][0] == _buffer_digest(ref)            # rank-count independent content
    assert out[0][1] == out[1][1] == ref_results
    assert out[0][2] and out[1][2]                                  # replicas identical after all-reduced steps


def test_arrays_roundtrip_single_process():
    """compact arrays -> the reference's (boards, actions, policies, qs, winner) tuples, exactly."""
    from alpha_zero_b200 import games
    from alpha_zero_b200.selfplay import games_from_arrays
    for G in (games.Gomoku, games.Connect4, games.TicTacToe):
        mine = [_fake_game(G, g, iters=30) for g in range(3)]
        spec, arr = _arrays_of(G, mine)
        back, vis = games_from_arrays(spec, arr, return_visits=True)
        for (boards, actions, policies, qs, winner), ((b0, a0, p0, q0, w0), v0), v in zip(back, mine, vis):
            assert winner == w0 and qs == q0 and [tuple(a) for a in a0] == actions
            assert all(x.dtype == np.float32 and np.array_equal(x, y) for x, y in zip(boards, b0))
            assert all(x.dtype == np.float64 and np.array_equal(x, y) for x, y in zip(policies, p0))
            assert np.array_equal(v, v0)
