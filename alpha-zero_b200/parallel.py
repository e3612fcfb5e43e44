"""Multi-GPU self-play: games are the unit of parallelism (the reference plays them one after
another with no shared state, train.py:64-67).  One process per GPU; game g belongs to rank
g % world; there is NO collective inside a search.  The only exchange is the all-gather of the
finished trajectories into every rank's replay buffer, in a compact form

    per position : side bitboards 2 x nw x u64, side to move u8, root visit counts u16[A], root Q f64
    per game     : global id i32, plies u16, winner i8

(116 B per 7x7 position instead of the 788 B of the reference's f32 board + f64 policy).  The
receiver rebuilds exactly what Game.self_play returns (boards float32 [C,H,W], policies
visits / sum(visits) as float64, gomoku.py:138-141) and stores it with train.save_data_to_buffer
(canonicalisation + augmentation, train.py:30-49) in GLOBAL game order, so the buffer content does
not depend on the number of ranks.  Gradients are all-reduced in train.train.
"""
import numpy as np
import torch
import torch.distributed as dist


def world():
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def shard_games(n_games, rank=None, world_size=None):
    """Global ids of the games this rank plays: g % world == rank."""
    if rank is None:
        rank, world_size = world()
    return list(range(rank, n_games, world_size))


def _nw(Game):
    return 1 if Game.rows * Game.cols <= 64 else 4


def pack_games(Game, games, global_ids, visits):
    """games: list of (boards, actions, policies, qs, winner) as play_games returns them;
    visits: per game, int array [plies][A] of root visit counts.  -> uint8 array."""
    nw, A = _nw(Game), Game.action_dim
    cells = Game.rows * Game.cols
    out = []
    for gid, (boards, actions, _policies, qs, winner), vis in zip(global_ids, games, visits):
        plies = len(boards)
        out.append(np.array([gid], np.int32).tobytes())
        out.append(np.array([plies], np.uint16).tobytes())
        out.append(np.array([winner], np.int8).tobytes())
        words = np.zeros((plies, 2, nw), np.uint64)
        for i, b in enumerate(boards):
            for side in range(2):
                idx = np.nonzero(b[side].reshape(-1) == 1.0)[0]
                for c in idx:
                    words[i, side, c >> 6] |= np.uint64(1) << np.uint64(c & 63)
        assert cells <= 64 * nw
        out.append(words.tobytes())
        out.append(np.asarray(vis, np.uint16).reshape(plies, A).tobytes())
        out.append(np.asarray(qs, np.float64).tobytes())
        acts = np.array([[-1, -1] if a is None else list(a) for a in actions], np.int16)   # plies + 1 entries (gomoku.py:129)
        out.append(np.array([len(acts)], np.uint16).tobytes())
        out.append(acts.tobytes())
    return np.frombuffer(b"".join(out), np.uint8).copy()


def unpack_games(Game, blob):
    """Inverse of pack_games -> list of (global id, (boards, actions, policies, qs, winner))."""
    nw, A, C = _nw(Game), Game.action_dim, Game.feature_dim
    rows, cols = Game.rows, Game.cols
    buf = np.asarray(blob, np.uint8).tobytes()
    off, games = 0, []

    def take(dtype, count):
        nonlocal off
        n = np.dtype(dtype).itemsize * count
        a = np.frombuffer(buf, dtype, count, off)
        off += n
        return a

    while off < len(buf):
        gid = int(take(np.int32, 1)[0])
        plies = int(take(np.uint16, 1)[0])
        winner = int(take(np.int8, 1)[0])
        words = take(np.uint64, plies * 2 * nw).reshape(plies, 2, nw)
        vis = take(np.uint16, plies * A).reshape(plies, A).astype(np.int64)
        qs = [float(q) for q in take(np.float64, plies)]
        na = int(take(np.uint16, 1)[0])
        acts = [tuple(int(v) for v in a) for a in take(np.int16, na * 2).reshape(na, 2)]
        boards, policies = [], []
        for i in range(plies):
            b = np.zeros((C, rows, cols), np.float32)
            for side in range(2):
                for w in range(nw):
                    v = int(words[i, side, w])
                    while v:
                        low = v & -v
                        c = w * 64 + low.bit_length() - 1
                        b[side, c // cols, c % cols] = 1.0
                        v ^= low
            if C == 3:                                  # plane 2 = side to move (connect4.py:55-70): ply i is played by i % 2
                b[2] = float(i % 2)
            boards.append(b)
            policies.append(vis[i] / vis[i].sum())      # utils.py:46-55 (float64)
        games.append((gid, (boards, acts, policies, qs, winner)))
    return games


def all_gather_bytes(blob, device=None):
    """Variable-length all-gather of uint8 arrays: sizes first, then the padded payload."""
    rank, ws = world()
    if ws == 1:
        return [np.asarray(blob, np.uint8)]
    if device is None:
        device = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend() == "nccl" else torch.device("cpu")
    n = torch.tensor([len(blob)], dtype=torch.int64, device=device)
    sizes = [torch.zeros_like(n) for _ in range(ws)]
    dist.all_gather(sizes, n)
    sizes = [int(s.item()) for s in sizes]
    cap = max(max(sizes), 1)
    mine = torch.zeros(cap, dtype=torch.uint8, device=device)
    mine[:len(blob)] = torch.from_numpy(np.asarray(blob, np.uint8)).to(device)
    parts = [torch.empty(cap, dtype=torch.uint8, device=device) for _ in range(ws)]
    dist.all_gather(parts, mine)
    return [p[:s].cpu().numpy() for p, s in zip(parts, sizes)]


def gather_games(Game, games, global_ids, visits):
    """Every rank receives every finished game, sorted by global id."""
    blobs = all_gather_bytes(pack_games(Game, games, global_ids, visits))
    merged = [g for blob in blobs for g in unpack_games(Game, blob)]
    merged.sort(key=lambda t: t[0])
    return merged


def store_games(Game, buffer, merged):
    """train.collect_data's bookkeeping (train.py:69-75) for gathered games, in global order."""
    from .train import save_data_to_buffer
    results = [0, 0, 0]
    for _gid, (boards, actions, policies, qs, winner) in merged:
        reward = 0 if winner == -1 else 1
        results[2 if winner == -1 else winner] += 1
        save_data_to_buffer(Game, buffer, (boards, actions, policies, qs, winner, reward))
    return results


def collect_data_distributed(Game, model, buffer, iterations, mcts_iter, vl=1, dirichlet=True):
    """train.collect_data over all ranks: this rank plays games rank, rank + world, ... on its GPU,
    then all ranks exchange trajectories and fill identical buffers."""
    from .train import play_games
    ids = shard_games(iterations)
    games, visits = play_games(Game, model, len(ids), mcts_iter, vl=vl, dirichlet=dirichlet, return_visits=True) if ids else ([], [])
    merged = gather_games(Game, games, ids, visits)
    return store_games(Game, buffer, merged)


def broadcast_model(model, src=0):
    """Same weights on every rank (after a load, or to guard against drift)."""
    _rank, ws = world()
    if ws == 1:
        return
    for p in model.state_dict().values():
        dist.broadcast(p, src)
