"""Data collection and the optimiser step with the reference's entry points (train.py).

collect_data plays `iterations` self-play games CONCURRENTLY on the GPU forest (the reference
plays them one after another, train.py:64-67; games are independent, so only the order in
which the global NumPy RNG is consumed differs).  save_data_to_buffer / rotate_data /
flip_data produce exactly the reference's samples (train.py:8-49).  train() is the
reference's Adam + soft-target CE + MSE + explicit L2 step (train.py:85-123), with an optional
NCCL gradient all-reduce when torch.distributed is initialised.
"""
import time
from concurrent.futures import ThreadPoolExecutor

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


def _moves_from_actions(Game, boards, acts):
    """(row, col) of action index `acts[i]` on boards[i]: get_action_idx inverted for a whole batch
    (games/connect4.py:27-41: the stone drops to the lowest empty row of the column)."""
    if Game.name == "connect4":
        n = np.arange(len(acts))
        empty = (boards[:, 0] + boards[:, 1]) == 0                    # [n, rows, cols]
        return empty[n, :, acts].sum(axis=1).astype(np.int64) - 1, acts.astype(np.int64)
    return (acts // Game.cols).astype(np.int64), (acts % Game.cols).astype(np.int64)


def _wins(Game, boards, movers, r, c):
    """check_winner for a batch: a run of >= win_length of the mover's stones through (r, c)
    (games/*: check_winner looks along the four lines through the last move only)."""
    n = np.arange(len(r))
    mine = boards[n, movers] == 1                                      # [n, rows, cols]
    won = np.zeros(len(r), bool)
    for dr, dc in ((0, 1), (1, 0), (1, 1), (1, -1)):
        run = np.ones(len(r), np.int64)
        for s in (1, -1):
            going = np.ones(len(r), bool)
            for k in range(1, Game.win_length):
                rr, cc = r + s * k * dr, c + s * k * dc
                ok = going & (rr >= 0) & (rr < Game.rows) & (cc >= 0) & (cc < Game.cols)
                ok[ok] = mine[n[ok], rr[ok], cc[ok]]
                run += ok
                going = ok
        won |= run >= Game.win_length
    return won


def play_games(Game, model, n_games, mcts_iter, vl=1, dirichlet=True, return_visits=False, concurrent=None, seed=None):
    """n_games self-play games, `concurrent` of them in flight at once on the GPU forest (default: all);
    a finished game's slot is refilled with the next game until n_games have been started.  Returns the
    reference's 5-tuple per game in game order (and, on request, the raw root visit counts per ply).

    With a `Net` / `B200Net` / `HashEvaluator` the whole loop runs on the device (selfplay.SelfPlayEngine: move choice,
    rules, refill and Dirichlet noise are kernels; the host reads one status word per ply).  Other callables (external
    evaluators) and vanilla mode (model=None) use the batched host loop below, one NumPy step per ply."""
    from .ai.mcts import HashEvaluator
    from .ai.nn import Net, B200Net
    if isinstance(model, (Net, B200Net, HashEvaluator)) and type(model).__name__ != "UCB1HashEvaluator":
        from .selfplay import play_games_device, games_from_arrays
        from .forest import game_spec
        arr = play_games_device(Game, model, n_games, mcts_iter, vl=vl, dirichlet=dirichlet, concurrent=concurrent, seed=seed,
                                return_arrays=True)
        return games_from_arrays(game_spec(Game.name, Game.rows), arr, return_visits=return_visits)
    return _play_games_host(Game, model, n_games, mcts_iter, vl, dirichlet, return_visits, concurrent)


def _play_games_host(Game, model, n_games, mcts_iter, vl=1, dirichlet=True, return_visits=False, concurrent=None):
    """Host-driven variant: per ply batched NumPy work -- pi = visits / sum (utils.py:46-55), a move drawn from pi
    while ply < sample_until (np.random.choice's inverse CDF, searchsorted side='right') else the most visited action,
    make_move, check_winner.  Ties of the most visited action resolve to the lowest action index here (the reference
    takes the first maximum in child order, which the device engine reproduces)."""
    C, R, W_, A = Game.feature_dim, Game.rows, Game.cols, Game.action_dim
    slots = min(n_games, concurrent or n_games)
    boards = np.zeros((slots, C, R, W_), np.float32)
    players = np.zeros(slots, np.int32)
    plies = np.zeros(slots, np.int32)
    alive = np.ones(slots, np.uint8)
    gid = np.arange(slots)                                             # game played in each slot
    started = slots
    winners = np.full(n_games, -1, np.int64)
    log = dict(gid=[], snap=[], pi=[], vis=[], q=[], r=[], c=[])       # one entry per ply: arrays over the live games
    sample_until = (8 if Game.name == "gomoku" else Game.state_dim + 1) if model else 0    # gomoku.py:143 / connect4.py:133
    # Dirichlet noise of the NEXT ply is drawn by a helper thread while the GPU searches (small-alpha gamma
    # sampling is the most expensive host step); its generator is seeded from the global RNG, so runs repeat.
    dirichlet = bool(dirichlet and model)                           # vanilla mode never mixes noise in (ai/mcts.py:42,52)
    noise_rng = np.random.default_rng(np.random.randint(0, 2 ** 31 - 1, size=4))
    pool = ThreadPoolExecutor(1) if dirichlet else None
    draw = (lambda: noise_rng.dirichlet([0.03] * A, size=slots)) if dirichlet else None
    pending = pool.submit(draw) if dirichlet else None
    while alive.any():
        noise = pending.result() if dirichlet else None
        if dirichlet:
            pending = pool.submit(draw)
        visits, q = MCTS.mcts_batch(model, boards, players, plies, Game, mcts_iter, noise=noise, vl=vl, active=alive)
        live = np.nonzero(alive)[0]
        vis = visits[live].astype(np.float64)
        pi = vis / vis.sum(axis=1, keepdims=True)
        # one uniform per game -> inverse-CDF draw from pi (np.random.choice's own method, vectorised)
        u = np.random.random_sample(len(live))
        cdf = np.cumsum(pi, axis=1)
        cdf /= cdf[:, -1:]
        drawn = np.minimum((cdf <= u[:, None]).sum(axis=1), A - 1)          # searchsorted(cdf, u, side='right')
        zero = visits[live][np.arange(len(live)), drawn] == 0              # u exactly on a CDF step: never an unvisited action
        drawn = np.where(zero, np.argmax(visits[live], axis=1), drawn)
        acts = np.where(plies[live] < sample_until, drawn, np.argmax(visits[live], axis=1))   # ties: lowest action index
        snap = boards[live].copy()
        r, c = _moves_from_actions(Game, snap, acts)
        movers = players[live].astype(np.int64)
        boards[live, movers, r, c] = 1
        if C == 3:
            boards[live, 2] = (1 - movers)[:, None, None]             # plane 2 = side to move next (connect4.py:55-70)
        players[live] = 1 - movers
        plies[live] += 1
        won = _wins(Game, boards[live], movers, r, c)
        over = won | (plies[live] == Game.state_dim)
        for k, v in (("gid", gid[live].copy()), ("snap", snap), ("pi", pi), ("vis", visits[live]), ("q", q[live]), ("r", r), ("c", c)):
            log[k].append(v)
        done = live[over]
        winners[gid[done]] = np.where(won[over], movers[over], -1)
        for s in done:                                                # refill the slot with the next game, or retire it
            if started < n_games:
                gid[s] = started
                started += 1
                boards[s] = 0
                players[s] = 0
                plies[s] = 0
            else:
                alive[s] = 0
    if pool is not None:
        pending.cancel()
        pool.shutdown(wait=True)
    # per-game records: a game lives in one slot through consecutive plies, so a stable sort by game id keeps ply order
    cat = {k: np.concatenate(v) for k, v in log.items()}
    order = np.argsort(cat["gid"], kind="stable")
    cuts = np.cumsum(np.bincount(cat["gid"], minlength=n_games))[:-1]
    part = {k: np.split(v[order], cuts) for k, v in cat.items()}
    games, all_visits = [], []
    for g in range(n_games):
        actions = [(-1, -1)] + list(zip(part["r"][g].tolist(), part["c"][g].tolist()))          # gomoku.py:128
        games.append((list(part["snap"][g]), actions, list(part["pi"][g]), part["q"][g].tolist(), int(winners[g])))
        all_visits.append(part["vis"][g])
    if return_visits:
        return games, all_visits
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


last_stats = {}        # filled by train(): iterations, world, gradient bytes, all-reduce time left exposed (bench.py's train leg)


class _PolicyValueLoss(torch.autograd.Function):
    """train.py:110-112 as ONE kernel (az_loss_policy_value, csrc/az_replay.cu): soft-target cross entropy + MSE, forward
    and backward together; returns (policy_loss, value_loss) and hands the stored gradients back in backward."""

    @staticmethod
    def forward(ctx, logits, values, pis, zs):
        from . import lib as _lib
        L = _lib.load()
        logits, values, pis, zs = (t.contiguous().float() for t in (logits, values, pis, zs))
        B, A = logits.shape
        dl, dv = torch.empty_like(logits), torch.empty_like(values)
        out = torch.empty(2, dtype=torch.float32, device=logits.device)
        _lib.check(L.az_loss_policy_value(logits.data_ptr(), values.data_ptr(), pis.data_ptr(), zs.data_ptr(), int(B), int(A), dl.data_ptr(),
                                          dv.data_ptr(), out.data_ptr(), torch.cuda.current_stream().cuda_stream), "az_loss_policy_value")
        ctx.save_for_backward(dl, dv)
        return out[0], out[1]

    @staticmethod
    def backward(ctx, g_policy, g_value):
        dl, dv = ctx.saved_tensors
        return dl * g_policy, dv * g_value, None, None


def policy_value_loss(logits, values, pis, zs):
    """(policy_loss, value_loss) of train.py:110-112; the fused CUDA kernel for CUDA tensors, the reference's torch ops on CPU."""
    if logits.is_cuda:
        return _PolicyValueLoss.apply(logits, values, pis, zs)
    return (torch.mean(torch.sum(-pis * torch.log_softmax(logits, dim=1), dim=1)), torch.nn.functional.mse_loss(values, zs))


class _GradBuckets:
    """Data-parallel gradient averaging overlapped with backward: every parameter's .grad is a view into one flat buffer
    cut into buckets (in reverse parameter order = the order backward produces them); when the last gradient of a bucket
    has been accumulated its all-reduce starts asynchronously (NCCL over NVLink) while backward continues."""

    def __init__(self, model, bucket_bytes=4 << 20):
        import torch.distributed as dist
        self.dist, self.world = dist, dist.get_world_size()
        params = [p for p in model.parameters() if p.requires_grad]
        self.flat = torch.zeros(sum(p.numel() for p in params), dtype=params[0].dtype, device=params[0].device)
        off = self.flat.numel()
        self.buckets, cur, cur_hi = [], [], off                     # bucket = [lo, hi, params, pending]
        for p in reversed(params):
            off -= p.numel()
            p.grad = self.flat[off:off + p.numel()].view_as(p)
            cur.append(p)
            if (cur_hi - off) * self.flat.element_size() >= bucket_bytes or off == 0:
                self.buckets.append(dict(lo=off, hi=cur_hi, params=cur, left=len(cur)))
                cur, cur_hi = [], off
        self.of = {id(p): b for b in self.buckets for p in b["params"]}
        self.work = []
        self.exposed = []                                 # CUDA event pairs around the waits of finish()
        self.hooks = [p.register_post_accumulate_grad_hook(self._ready) for p in params]

    def _ready(self, p):
        b = self.of[id(p)]
        b["left"] -= 1
        if b["left"] == 0:
            self.work.append(self.dist.all_reduce(self.flat[b["lo"]:b["hi"]], async_op=True))

    def zero(self):
        self.flat.zero_()
        for b in self.buckets:
            b["left"] = len(b["params"])

    def finish(self):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        timed = self.flat.is_cuda
        if timed:
            a.record()
        for w in self.work:
            w.wait()
        if timed:
            b.record()                                   # what of the all-reduces backward did not hide
            self.exposed.append((a, b))
        self.work.clear()
        self.flat /= self.world

    def close(self):
        for h in self.hooks:
            h.remove()


def train(model, batch_size, buffer, train_iterations, lr, device):
    """train.py:85-123: fresh Adam, loss = soft-target CE + MSE + 1e-4 * sum(w^2) over every
    parameter whose name has no 'bias' (the 'LayerNorm' filter of train.py:106 never matches,
    so LayerNorm weights, cls_token and pos_embedding ARE regularised -- kept as is).
    `buffer` is a ReplayBuffer (host deque, one H2D per step as in the reference, train.py:98-99) or a
    DeviceReplayBuffer (samples are CUDA tensors already).  Under torch.distributed (world > 1) every rank samples its own
    batch and the gradients are averaged with bucketed all-reduces overlapped with backward."""
    import torch.distributed as dist
    model.to(device)
    model.train()
    opt = torch.optim.Adam(model.parameters(), lr=lr)
    ddp = dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1
    buckets = _GradBuckets(model) if ddp else None
    reg = [p for name, p in model.named_parameters() if "bias" not in name and "LayerNorm" not in name]
    loss = policy_loss = value_loss = l2 = torch.zeros(())
    t0 = time.time()
    for i in range(train_iterations):
        states, pis, zs = (t.to(device, non_blocking=True) for t in buffer.sample(batch_size))
        l2 = torch.stack([torch.sum(p * p) for p in reg]).sum()
        logits, values = model(states)
        policy_loss, value_loss = policy_value_loss(logits, values, pis, zs)
        loss = policy_loss + value_loss + 1e-4 * l2
        if buckets:
            buckets.zero()
        else:
            opt.zero_grad()
        loss.backward()
        if buckets:                                     # data-parallel ranks: average gradients over NVLink
            buckets.finish()
        opt.step()
        if (i + 1) % 50 == 0:
            now = time.time()
            print("train_iter:", i + 1, "loss:", loss.item(), "policy_loss:", policy_loss.item(), "value_loss:",
                  value_loss.item(), "l2_loss:", l2.item(), "time:", now - t0)
            t0 = now
    last_stats.clear()
    last_stats.update(iterations=train_iterations, batch_size=batch_size, world=dist.get_world_size() if ddp else 1,
                      grad_bytes=sum(p.numel() * p.element_size() for p in model.parameters()))
    if buckets:
        if buckets.flat.is_cuda:
            torch.cuda.synchronize()
        last_stats.update(allreduce_exposed_ms=sum(a.elapsed_time(b) for a, b in buckets.exposed), buckets=len(buckets.buckets))
        buckets.close()
        for p in model.parameters():                    # leave ordinary, independent .grad tensors behind
            p.grad = None
    return loss.item(), policy_loss.item(), value_loss.item(), l2.item()
