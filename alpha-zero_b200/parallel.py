"""Multi-GPU self-play: games are the unit of parallelism (the reference plays them one after
another with no shared state, train.py:64-67).  One process per GPU; game g belongs to rank
g % world (local game k of rank r has global id r + k * world); there is NO collective inside a
search or a game.  The only exchange is the all-gather of the finished trajectories, straight from
the device-resident trajectory store of the self-play engine (csrc/az_selfplay.cu) in compact form

    per game     : plies i32, winner i32, first recorded ply i32
    per position : side bitboards 2 x nw x u64, root visit counts u16[A], root Q f64, move (cell) i16

(124 B per 7x7 position instead of the 788 B of the reference's f32 board + f64 policy), ONE
all_gather_into_tensor of the byte image of those arrays over NCCL / NVLink.  Because every random
draw of a game is keyed by its GLOBAL id (Philox, az_selfplay.cu) and games are merged in global
order, the gathered trajectories -- and the replay buffer filled from them -- do not depend on
the number of ranks (given a deterministic evaluator).  Gradients are all-reduced in train.train.
"""
import numpy as np
import torch
import torch.distributed as dist

TRAJ_KEYS = ("plies", "winner", "first", "words", "visits", "q", "action")


def world():
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def shard_games(n_games, rank=None, world_size=None):
    """Global ids of the games this rank plays: g % world == rank."""
    if rank is None:
        rank, world_size = world()
    return list(range(rank, n_games, world_size))


def _as_tensor(x):
    if isinstance(x, torch.Tensor):
        return x
    a = np.ascontiguousarray(x)
    if a.dtype == np.uint64:
        a = a.view(np.int64)
    elif a.dtype == np.uint16:
        a = a.view(np.int16)
    return torch.from_numpy(a)


def gather_trajectories(local, n_total):
    """local: dict of TRAJ_KEYS -> tensor (CUDA for NCCL, CPU for gloo) or array with leading dim = this rank's games
    (ids rank, rank + world, ...).  Returns the same dict for ALL n_total games in global id order, on the same device.
    One collective: the arrays' bytes are concatenated, padded to ceil(n_total / world) games and all-gathered."""
    rank, ws = world()
    t = {k: _as_tensor(local[k]) for k in TRAJ_KEYS}
    if ws == 1:
        return {k: v[:n_total] for k, v in t.items()}
    gmax = (n_total + ws - 1) // ws
    dev = t["q"].device
    parts, meta = [], []
    for k in TRAJ_KEYS:
        x = t[k].contiguous()
        if x.shape[0] < gmax:                                          # ranks with one game fewer: zero padding (plies = 0)
            pad = torch.zeros((gmax - x.shape[0],) + tuple(x.shape[1:]), dtype=x.dtype, device=dev)
            x = torch.cat([x, pad], 0)
        meta.append((k, x.dtype, tuple(x.shape[1:])))
        parts.append(x[:gmax].reshape(-1).view(torch.uint8))
    mine = torch.cat(parts)
    flat = torch.empty(ws * mine.numel(), dtype=torch.uint8, device=dev)
    dist.all_gather_into_tensor(flat, mine)
    out = flat.view(ws, mine.numel())
    merged, off = {}, 0
    for k, dt, tail in meta:
        nbytes = gmax * int(np.prod(tail, dtype=np.int64)) * torch.empty((), dtype=dt).element_size()
        x = out[:, off:off + nbytes].contiguous().view(dt).reshape((ws, gmax) + tail)
        merged[k] = x.transpose(0, 1).reshape((ws * gmax,) + tail)[:n_total]     # [k][rank] -> global id rank + k * world
        off += nbytes
    return merged


def trajectories_to_host(t):
    """dict of tensors -> dict of NumPy arrays with the engine's dtypes (u64 words, u16 visits)."""
    out = {k: v.detach().cpu().numpy() for k, v in t.items() if k in TRAJ_KEYS}
    out["words"] = out["words"].view(np.uint64)
    out["visits"] = out["visits"].view(np.uint16)
    return out


def store_games(Game, buffer, games):
    """train.collect_data's bookkeeping (train.py:69-75) for games in global order."""
    from .train import save_data_to_buffer
    results = [0, 0, 0]
    for boards, actions, policies, qs, winner in games:
        reward = 0 if winner == -1 else 1
        results[2 if winner == -1 else winner] += 1
        save_data_to_buffer(Game, buffer, (boards, actions, policies, qs, winner, reward))
    return results


def collect_data_distributed(Game, model, buffer, iterations, mcts_iter, vl=1, dirichlet=True, seed=None, concurrent=None):
    """train.collect_data over all ranks: this rank plays games rank, rank + world, ... on its GPU (device-side
    self-play), all ranks exchange the compact trajectories over NCCL and fill identical buffers in global game order.
    `seed` must be the same on every rank (default: drawn from rank 0's NumPy RNG and broadcast)."""
    from .forest import game_spec
    from .selfplay import play_games_device, games_from_arrays
    rank, ws = world()
    if seed is None:
        s = torch.tensor([int(np.random.randint(0, 2 ** 31 - 1)), int(np.random.randint(0, 2 ** 31 - 1))], dtype=torch.int64,
                         device="cuda" if ws > 1 and dist.get_backend() == "nccl" else "cpu")
        if ws > 1:
            dist.broadcast(s, 0)
        seed = int(s[0].item()) * 2654435761 + int(s[1].item())
    ids = shard_games(iterations, rank, ws)
    g = game_spec(Game.name, Game.rows)
    if ids:
        sp = play_games_device(Game, model, len(ids), mcts_iter, vl=vl, dirichlet=dirichlet, concurrent=concurrent, seed=seed,
                               id_first=rank, id_stride=ws, return_engine=True)
        local = sp.device_tensors()
    else:                                                               # more ranks than games
        local, sp = empty_trajectories(g, 0, "cuda"), None
    merged = gather_trajectories(local, iterations)
    games = games_from_arrays(g, trajectories_to_host(merged))
    if sp is not None:
        sp.close()
    return store_games(Game, buffer, games)


def collect_data_device(Game, model, dbuffer, iterations, mcts_iter, vl=1, dirichlet=True, seed=0, concurrent=None):
    """train.collect_data with everything resident in HBM: this rank plays its shard of `iterations` games on the device,
    the compact trajectories are all-gathered over NCCL and appended to a DeviceReplayBuffer (canonicalisation, z and the
    dihedral variants run as one kernel) in global game order -- no host copy of a position anywhere.  Every rank ends
    with the same buffer.  Returns [wins of side 0, wins of side 1, draws] as train.collect_data does (train.py:69-73)."""
    from .forest import game_spec
    from .selfplay import play_games_device
    rank, ws = world()
    ids = shard_games(iterations, rank, ws)
    g = game_spec(Game.name, Game.rows)
    sp = None
    if ids:
        sp = play_games_device(Game, model, len(ids), mcts_iter, vl=vl, dirichlet=dirichlet, concurrent=concurrent, seed=seed,
                               id_first=rank, id_stride=ws, return_engine=True)
        local = sp.device_tensors()
    else:
        local = empty_trajectories(g, 0, "cuda")
    merged = gather_trajectories(local, iterations)
    dbuffer.add_trajectories(merged)
    w = merged["winner"]
    results = [int((w == 0).sum().item()), int((w == 1).sum().item()), int((w == -1).sum().item())]
    if sp is not None:
        sp.close()
    return results


def empty_trajectories(g, n, device="cpu"):
    P, A = g.state_dim, g.action_dim
    z = lambda shape, dt: torch.zeros(shape, dtype=dt, device=device)
    return dict(plies=z((n,), torch.int32), winner=z((n,), torch.int32), first=z((n,), torch.int32), words=z((n, P, 2, g.nw), torch.int64),
                visits=z((n, P, A), torch.int16), q=z((n, P), torch.float64), action=z((n, P), torch.int16))


def broadcast_model(model, src=0):
    """Same weights on every rank (after a load, or to guard against drift)."""
    _rank, ws = world()
    if ws == 1:
        return
    for p in model.state_dict().values():
        dist.broadcast(p, src)
