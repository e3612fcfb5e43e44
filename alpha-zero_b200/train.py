"""Data collection and the optimiser step with the reference's entry points (train.py).

collect_data plays `iterations` self-play games CONCURRENTLY on the GPU forest (the reference
plays them one after another, train.py:64-67; games are independent, so only the order in
which the global NumPy RNG is consumed differs).  save_data_to_buffer / rotate_data /
flip_data produce exactly the reference's samples (train.py:8-49).  train() is the
reference's Adam + soft-target CE + MSE + explicit L2 step (train.py:85-123), with an optional
NCCL gradient all-reduce when torch.distributed is initialised.
"""
import time

import numpy as np
import torch

from .ai import MCTS


def rotate_data(board, policy, k):
    rows, cols = board.shape[1:]
    return np.rot90(board, k=k, axes=(1, 2)).copy(), np.rot90(policy.reshape(rows, cols), k=k).flatten()


def flip_data(board, policy, mode):
    rows, cols = board.shape[1:]
    p2 = policy.reshape(rows, cols)
    if mode == "lr":
        return np.flip(board, axis=2).copy(), np.flip(p2, axis=1).flatten()
    return np.flip(board, axis=1).copy(), np.flip(p2, axis=0).flatten()


def save_data_to_buffer(Game, buffer, data):
    """train.py:30-49.  Ply i was played by side i%2; z = +-reward; canonical board; plies 0 and
    1 stored once, later plies in the 8 dihedral variants rot0, rot0-lr, rot0-tb, rot90,
    rot90-lr, rot90-tb, rot180, rot270.  Boards whose policy is not a rows x cols grid
    (Connect4) get identity + left-right mirror instead -- a build extension, the reference
    cannot augment them at all."""
    boards, _actions, policies, _qs, winner, reward = data
    side = 0
    for i, (board, policy) in enumerate(zip(boards, policies)):
        z = [reward] if side == winner else [-reward]
        canon = Game.get_canonical_board(board, side)
        side = 1 - side
        if i < 2:
            buffer.add(canon, policy, z)
            continue
        if Game.action_dim != Game.rows * Game.cols or Game.rows != Game.cols:
            buffer.add(canon, policy, z)
            buffer.add(np.flip(canon, axis=2).copy(), policy[::-1].copy(), z)
            continue
        for k in range(4):
            b_rot, p_rot = rotate_data(canon, policy, k)
            buffer.add(b_rot, p_rot, z)
            if k < 2:
                buffer.add(*flip_data(b_rot, p_rot, "lr"), z)
                buffer.add(*flip_data(b_rot, p_rot, "tb"), z)


def play_games(Game, model, n_games, mcts_iter, vl=1, dirichlet=True, return_visits=False):
    """n_games concurrent self-play games; returns the reference's 5-tuple per game (and, on request,
    the raw root visit counts per ply -- the compact form parallel.pack_games ships between ranks)."""
    boards = np.stack([Game().board for _ in range(n_games)])
    players = np.zeros(n_games, np.int32)
    plies = np.zeros(n_games, np.int32)
    alive = np.ones(n_games, np.uint8)
    rec = [dict(boards=[], actions=[(-1, -1)], policies=[], qs=[], winner=-1, visits=[]) for _ in range(n_games)]
    sample_until = 8 if Game.name == "gomoku" else Game.state_dim + 1          # gomoku.py:143 / connect4.py:133
    A = Game.action_dim
    while alive.any():
        noise = np.random.dirichlet([0.03] * A, size=n_games) if dirichlet else None
        visits, q = MCTS.mcts_batch(model, boards, players, plies, Game, mcts_iter, noise=noise, vl=vl, active=alive)
        for g in np.nonzero(alive)[0]:
            pi = visits[g] / visits[g].sum()
            r = rec[g]
            r["policies"].append(pi)
            r["visits"].append(visits[g].copy())
            r["boards"].append(boards[g].copy())
            r["qs"].append(q[g])
            legal = {Game.get_action_idx(m): m for m in Game.get_valid_moves(boards[g])}
            if plies[g] < sample_until:
                a = int(np.random.choice(A, p=pi))
            else:
                a = int(np.argmax(visits[g]))           # ties: lowest action index (node.py:81 keeps child order)
            move = legal[a]
            mover = int(players[g])
            players[g] = Game.make_move(boards[g], mover, move)
            plies[g] += 1
            r["actions"].append(move)
            w = Game.check_winner(boards[g], mover, move)
            if w != -1 or plies[g] == Game.state_dim:
                r["winner"] = w
                alive[g] = 0
    games = [(r["boards"], r["actions"], r["policies"], r["qs"], r["winner"]) for r in rec]
    if return_visits:
        return games, [np.stack(r["visits"]) for r in rec]
    return games


def collect_data(Game, model, buffer, iterations, mcts_iter, display=False, vl=1):
    """train.py:54-83 with all `iterations` games in flight at once."""
    if hasattr(model, "eval"):
        model.eval()
    results = [0, 0, 0]
    t0 = time.time()
    games = play_games(Game, model, iterations, mcts_iter, vl=vl)
    for boards, actions, policies, qs, winner in games:
        reward = 0 if winner == -1 else 1
        results[2 if winner == -1 else winner] += 1
        save_data_to_buffer(Game, buffer, (boards, actions, policies, qs, winner, reward))
    dt = time.time() - t0
    Game.logger.debug(f"collect_data iter:{iterations} total time: {dt}s average time per game: "
                      f"{dt / max(iterations, 1)}s\ngame results: {results}")
    print(f"buffer status:{buffer.size()}")
    return results


def train(model, batch_size, buffer, train_iterations, lr, device):
    """train.py:85-123: fresh Adam, loss = soft-target CE + MSE + 1e-4 * sum(w^2) over every
    parameter whose name has no 'bias' (the 'LayerNorm' filter of train.py:106 never matches,
    so LayerNorm weights, cls_token and pos_embedding ARE regularised -- kept as is)."""
    import torch.distributed as dist
    model.to(device)
    model.train()
    opt = torch.optim.Adam(model.parameters(), lr=lr)
    ddp = dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1
    loss = policy_loss = value_loss = l2 = torch.zeros(())
    t0 = time.time()
    for i in range(train_iterations):
        states, pis, zs = (t.to(device) for t in buffer.sample(batch_size))
        l2 = sum(torch.sum(p ** 2) for name, p in model.named_parameters() if "bias" not in name and "LayerNorm" not in name)
        logits, values = model(states)
        policy_loss = torch.mean(torch.sum(-pis * torch.log_softmax(logits, dim=1), dim=1))
        value_loss = torch.nn.functional.mse_loss(values, zs)
        loss = policy_loss + value_loss + 1e-4 * l2
        opt.zero_grad()
        loss.backward()
        if ddp:                                         # data-parallel ranks: average gradients over NVLink
            flat = torch.cat([p.grad.reshape(-1) for p in model.parameters()])
            dist.all_reduce(flat)
            flat /= dist.get_world_size()
            off = 0
            for p in model.parameters():
                p.grad.copy_(flat[off:off + p.numel()].view_as(p))
                off += p.numel()
        opt.step()
        if (i + 1) % 50 == 0:
            now = time.time()
            print("train_iter:", i + 1, "loss:", loss.item(), "policy_loss:", policy_loss.item(), "value_loss:",
                  value_loss.item(), "l2_loss:", l2.item(), "time:", now - t0)
            t0 = now
    return loss.item(), policy_loss.item(), value_loss.item(), l2.item()
