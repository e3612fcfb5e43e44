"""CPU oracle for the self-play hot path -- TEST INFRASTRUCTURE, NOT PRODUCT.

A plain-Python/numpy restatement of the reference's algorithm (ehgus8/alpha-zero):
rules of the three games, the PUCT search of ``ai/mcts.py`` + ``ai/node.py`` +
``utils.py`` and the trajectory -> replay-buffer transform of ``train.py``.  Every
function cites the reference ``file:line`` it follows.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl reference`` leg may
import this module, and only as the checker or the timed CPU baseline; the product
package (``alpha-zero_b200/``) never imports it.

Pinning: the reference ships no tests or golden vectors, so this
oracle is pinned against OUTPUTS OF THE REFERENCE ITSELF, generated in the build
container by ``oracle/gen_golden.py`` (which imports a scratch copy of /root/reference)
and committed under ``tests/golden/``.  ``tests/test_oracle_golden.py`` replays them.

Numeric model: the reference runs under NumPy >= 2 (NEP 50 promotion).  ``Node.prior``
is an ``np.float32`` scalar for ordinary nodes, an ``np.float64`` for root children when
Dirichlet noise is mixed in; ``Node.value`` is a Python float (f64).  We keep the very
same scalar types so that Python/NumPy performs the very same roundings.
"""
from __future__ import annotations

import math
from collections import deque

import numpy as np

MASK64 = (1 << 64) - 1


# --------------------------------------------------------------------------------------
# Game descriptors.  The reference uses classes as namespaces (games/game.py:4-38); the
# oracle uses one small record per game plus free functions taking it.
# --------------------------------------------------------------------------------------
class GameSpec:
    __slots__ = ("name", "gid", "rows", "cols", "planes", "k", "action_dim", "state_dim")

    def __init__(self, name, gid, rows, cols, planes, k, action_dim):
        self.name, self.gid, self.rows, self.cols = name, gid, rows, cols
        self.planes, self.k, self.action_dim = planes, k, action_dim
        self.state_dim = rows * cols

    def __repr__(self):
        return f"GameSpec({self.name} {self.rows}x{self.cols} k={self.k})"


TICTACTOE = GameSpec("tictactoe", 0, 3, 3, 3, 3, 9)       # games/tictactoe.py:9-18
CONNECT4 = GameSpec("connect4", 1, 6, 7, 3, 4, 7)         # games/connect4.py:7-14
GOMOKU = GameSpec("gomoku", 2, 7, 7, 2, 5, 49)            # games/gomoku.py:9-17


def gomoku_spec(n):
    """Gomoku with the class attributes patched to n x n."""
    return GameSpec("gomoku", 2, n, n, 2, 5, n * n)


SPECS = {"tictactoe": TICTACTOE, "connect4": CONNECT4, "gomoku": GOMOKU}

_DIRS = ((0, 1), (1, 0), (1, 1), (1, -1))                 # gomoku.py:87,99 ; tictactoe.py:75


def new_board(g):
    return np.zeros((g.planes, g.rows, g.cols), dtype=np.float32)


def action_idx(g, a):
    """tictactoe.py:28-34, gomoku.py:42-48 (r*rows+c); connect4.py:27-29 (column)."""
    if g.gid == 1:
        return a[1]
    return a[0] * g.rows + a[1]


def make_move(g, board, player, a):
    """tictactoe.py:36-46, connect4.py:55-63, gomoku.py:50-58.  Returns next player."""
    r, c = a
    if g.gid == 1:
        if r is None:
            return player
        board[player, r, c] = 1
        board[2, :, :] = 1 - player
        return 1 - player
    if board[0, r, c] == 0 and board[1, r, c] == 0:
        board[player, r, c] = 1
        if g.planes == 3:
            board[2, :, :] = 1 - player
        return 1 - player
    return player                                          # occupied: no-op (reference prints)


def undo_move(g, board, cur, a):
    """tictactoe.py:48-51, connect4.py:66-70, gomoku.py:60-63."""
    r, c = a
    board[1 - cur, r, c] = 0
    if g.planes == 3:
        board[2, :, :] = 1 - cur


def run_length(g, board, player, a, dr, dc):
    """Stones of `player` in the line through `a` along +-(dr,dc), `a` itself counted as 1
    (the reference's dfs, e.g. gomoku.py:67-84, is exactly a contiguous-run count)."""
    r0, c0 = a
    n = 1
    for s in (1, -1):
        r, c = r0 + s * dr, c0 + s * dc
        while 0 <= r < g.rows and 0 <= c < g.cols and board[player, r, c] == 1:
            n += 1
            r += s * dr
            c += s * dc
    return n


def check_winner(g, board, player, a):
    """tictactoe.py:53-79, connect4.py:72-98, gomoku.py:65-91: `player` if a run >= k
    passes through the last move, else -1.  Overlines win."""
    for dr, dc in _DIRS:
        if run_length(g, board, player, a, dr, dc) >= g.k:
            return player
    return -1


def valid_moves(g, board):
    """Child order of the search.  tictactoe.py:81-83 (row-major empties);
    connect4.py:43-53 (columns ascending, (drop_row, col)); gomoku.py:93-106 (empty
    8-neighbours of stones collected into a Python set, returned as list(set))."""
    if g.gid == 0:
        return [(r, c) for r in range(g.rows) for c in range(g.cols)
                if board[0, r, c] == 0 and board[1, r, c] == 0]
    if g.gid == 1:
        out = []
        for c in range(g.cols):
            if board[0, 0, c] == 0 and board[1, 0, c] == 0:
                for r in range(g.rows - 1, -1, -1):       # connect4.py:31-41
                    if board[0, r, c] == 0 and board[1, r, c] == 0:
                        out.append((r, c))
                        break
        return out
    return [divmod(i, g.cols) for i in gomoku_valid_cells(g, board)]


def gomoku_insertion_sequence(g, board):
    """The exact sequence of .add() calls of gomoku.py:95-104 as cell indices
    (duplicates included): stones row-major; per stone dirs (0,1),(1,0),(1,1),(1,-1);
    per dir +dir first, then -dir; only empty in-bounds cells."""
    occ = (board[0] == 1) | (board[1] == 1)
    emp = (board[0] == 0) & (board[1] == 0)
    seq = []
    for r in range(g.rows):
        for c in range(g.cols):
            if occ[r, c]:
                for dr, dc in _DIRS:
                    for s in (1, -1):
                        rr, cc = r + s * dr, c + s * dc
                        if 0 <= rr < g.rows and 0 <= cc < g.cols and emp[rr, cc]:
                            seq.append(rr * g.cols + cc)
    return seq


# ---- CPython set emulation (Objects/setobject.c of CPython 3.12, the interpreter the
# reference's list(set) order comes from; restated from the published algorithm) -------
_XXP1, _XXP2, _XXP5 = 11400714785074694791, 14029467366897019727, 2870177450012600261


def cpython_tuple2_hash(r, c):
    """hash((r, c)) for small non-negative ints (Objects/tupleobject.c tuplehash,
    xxHash-style).  Returned as the unsigned 64-bit pattern."""
    acc = _XXP5
    for lane in (r, c):
        acc = (acc + lane * _XXP2) & MASK64
        acc = ((acc << 31) | (acc >> 33)) & MASK64
        acc = (acc * _XXP1) & MASK64
    acc = (acc + (2 ^ (_XXP5 ^ 3527539))) & MASK64
    if acc == MASK64:
        acc = 1546275796
    return acc


def cpython_set_order(hashes, seq):
    """Iteration order of a CPython set after `for x in seq: s.add(x)` starting empty.
    `hashes[x]` is the unsigned 64-bit hash of element x.  No deletions ever happen, so
    there are no dummy entries.  Returns the elements in slot order."""
    LINEAR = 9
    mask = 7
    table = [-1] * 8
    fill = 0

    def probe_insert(tab, msk, x, check_dup):
        h = hashes[x]
        perturb = h
        i = h & msk
        while True:
            lim = LINEAR if i + LINEAR <= msk else 0
            for j in range(lim + 1):
                e = tab[i + j]
                if e == -1:
                    tab[i + j] = x
                    return True
                if check_dup and e == x:
                    return False
            perturb >>= 5
            i = (i * 5 + 1 + perturb) & msk

    for x in seq:
        if probe_insert(table, mask, x, True):
            fill += 1
            if fill * 5 >= mask * 3:                      # set_add_entry resize rule
                minused = fill * 4 if fill <= 50000 else fill * 2
                newsize = 8
                while newsize <= minused:
                    newsize <<= 1
                old = table
                table = [-1] * newsize
                mask = newsize - 1
                for e in old:                            # set_table_resize: old slot order
                    if e != -1:
                        probe_insert(table, mask, e, False)
    return [e for e in table if e != -1]


_HASH_CACHE = {}


def gomoku_valid_cells(g, board):
    """Legal cell indices in reference child order, through the explicit set emulation
    (this is what the C oracle and the CUDA kernels restate)."""
    key = (g.rows, g.cols)
    hs = _HASH_CACHE.get(key)
    if hs is None:
        hs = [cpython_tuple2_hash(i // g.cols, i % g.cols) for i in range(g.rows * g.cols)]
        _HASH_CACHE[key] = hs
    seq = gomoku_insertion_sequence(g, board)
    if not seq:
        return [(g.rows // 2) * g.cols + g.cols // 2]     # gomoku.py:104-105
    return cpython_set_order(hs, seq)


def valid_moves_pyset(g, board):
    """Gomoku legal moves through a REAL Python set, statement by statement as
    gomoku.py:93-106 does; used to cross-check the emulation above."""
    s = set()
    for r in range(g.rows):
        for c in range(g.cols):
            if board[0, r, c] == 1 or board[1, r, c] == 1:
                for dr, dc in _DIRS:
                    if 0 <= r + dr < g.rows and 0 <= c + dc < g.cols and \
                            board[0, r + dr, c + dc] == 0 and board[1, r + dr, c + dc] == 0:
                        s.add((r + dr, c + dc))
                    if 0 <= r - dr < g.rows and 0 <= c - dc < g.cols and \
                            board[0, r - dr, c - dc] == 0 and board[1, r - dr, c - dc] == 0:
                        s.add((r - dr, c - dc))
    if len(s) == 0:
        s.add((g.rows // 2, g.cols // 2))
    return list(s)


def canonical_board(g, board, player):
    """gomoku.py:34-40: player 0 -> the board itself, player 1 -> planes 0/1 swapped.
    3-plane games: the reference has no live version; the shim is the author's legacy
    code in the comments of ai/mcts.py:126-137 (swap 0/1, keep plane 2)."""
    if player == 0:
        return board
    out = np.empty_like(board)
    out[0], out[1] = board[1], board[0]
    if g.planes == 3:
        out[2] = board[2]
    return out


# --------------------------------------------------------------------------------------
# Deterministic "hash evaluator": a stand-in policy/value function whose outputs are
# exact in fp32 on any IEEE machine (integer hash -> scaled by powers of two), so the CPU
# oracle, the unmodified reference (through its MCTS.cache hook, ai/mcts.py:40-41) and the
# CUDA path can be driven by IDENTICAL priors/values.  Build-defined; not in the reference.
# --------------------------------------------------------------------------------------
def splitmix64(x):
    x = (x + 0x9E3779B97F4A7C15) & MASK64
    z = x
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & MASK64
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & MASK64
    return z ^ (z >> 31)


def board_words(g, plane):
    """Occupancy of one f32 plane as 64-bit words, bit (r*cols+c)%64 of word (r*cols+c)//64."""
    flat = (np.asarray(plane).reshape(-1) == 1)
    nw = (g.rows * g.cols + 63) // 64
    out = []
    for w in range(nw):
        v = 0
        for i in np.nonzero(flat[w * 64:(w + 1) * 64])[0]:
            v |= 1 << int(i)
        out.append(v)
    return out


def ceil_log2(n):
    k = 0
    while (1 << k) < n:
        k += 1
    return k


def hash_eval_key(g, canon, seed):
    """64-bit key of a CANONICAL board (plane 0 = side to move's stones)."""
    h = splitmix64(seed & MASK64)
    for w in board_words(g, canon[0]):
        h = splitmix64(h ^ w)
    h = splitmix64(h ^ 0xA5A5A5A5A5A5A5A5)
    for w in board_words(g, canon[1]):
        h = splitmix64(h ^ w)
    if g.planes == 3:                                     # plane 2 = side to move, kept as is
        h = splitmix64(h ^ (1 if canon[2, 0, 0] == 1 else 0))
    return h


def hash_eval_from_key(g, h, qbits=16):
    """-> (priors f32[A] with sum <= 1, value f32 in [-1,1)).  All values are exact."""
    A = g.action_dim
    qm = (1 << qbits) - 1
    scale = np.float32(2.0 ** -(qbits + ceil_log2(A)))
    pri = np.empty(A, dtype=np.float32)
    for a in range(A):
        x = splitmix64((h + a + 1) & MASK64)
        pri[a] = np.float32(((x >> 40) & qm) + 1) * scale
    x = splitmix64((h + 0xFFFF) & MASK64)
    val = np.float32((int((x >> 40) & 0xFFFF) - 32768)) * np.float32(2.0 ** -15)
    return pri, val


class HashEvaluator:
    """evaluate(canon_board, player) -> (priors f32[A], value f32)."""

    def __init__(self, g, seed=0, qbits=16):
        self.g, self.seed, self.qbits = g, seed, qbits
        self.calls = 0

    def __call__(self, canon):
        self.calls += 1
        return hash_eval_from_key(self.g, hash_eval_key(self.g, canon, self.seed), self.qbits)


class TableEvaluator:
    """Replays logged (canonical board bytes -> priors, value) pairs, i.e. the reference's
    `MCTS.cache` preload route of SURVEY.md section 8c, for driving the oracle with the numbers
    a GPU net produced."""

    def __init__(self, table):
        self.table = table

    def __call__(self, canon):
        return self.table[np.ascontiguousarray(canon).tobytes()]


class TorchEvaluator:
    """The reference's own evaluation glue, ai/mcts.py:46-49: batch-1 forward on CPU,
    then exp(l)/sum(exp(l)) in float32 with NO max subtraction, NO legal masking."""

    def __init__(self, model):
        import torch
        self.torch, self.model = torch, model

    def __call__(self, canon):
        with self.torch.no_grad():
            logits, value = self.model(self.torch.from_numpy(np.ascontiguousarray(canon)).unsqueeze(0))
        l = logits.squeeze(0).numpy()
        return np.exp(l) / np.sum(np.exp(l)), np.float32(value.item())


# --------------------------------------------------------------------------------------
# Search tree.  ai/node.py:20-40.
# --------------------------------------------------------------------------------------
class Node:
    __slots__ = ("parent", "visit", "value", "ucb", "prior", "move_count", "action",
                 "player", "children")

    def __init__(self, parent, action, player, move_count, prior=0.0):
        self.parent = parent
        self.visit = 0
        self.value = 0            # W: f64 running sum, viewpoint of the player who moved in
        self.ucb = np.inf
        self.prior = prior
        self.move_count = move_count
        self.action = action      # reference name: prevAction
        self.player = player      # reference name: currentPlayer (side to move at node)
        self.children = []


def puct_scores(node):
    """utils.py:29-44, 'network' branch, c_puct = 1, no first-play urgency.  NumPy>=2
    scalar promotion makes this fp32 when prior is np.float32:
    s=(float)sqrt((double)Np); t=P*s; u=t/(float)(n+1); q=(float)(W/n); ucb=q+u."""
    root_n = node.visit
    for ch in node.children:
        if ch.visit == 0:
            ch.ucb = ch.prior * math.sqrt(root_n) / (ch.visit + 1)
        else:
            ch.ucb = (ch.value / ch.visit) + ch.prior * math.sqrt(root_n) / (ch.visit + 1)


def select_child(node):
    """node.py:42-47: first maximum in child order (max() keeps the earliest on ties)."""
    puct_scores(node)
    best = node.children[0]
    for ch in node.children[1:]:
        if ch.ucb > best.ucb:
            best = ch
    return best


def ucb1_scores(node):
    """utils.py:29-44, 'normal' branch (model=None): Python floats (f64) throughout,
    unvisited child: sqrt(2 ln N_p / 1) -- finite, not +inf; visited: W/n + sqrt(2 ln N_p / (n+1))."""
    for ch in node.children:
        if ch.visit == 0:
            ch.ucb = math.sqrt(2 * math.log(node.visit) / (ch.visit + 1))
        else:
            ch.ucb = (ch.value / ch.visit) + math.sqrt(2 * math.log(node.visit) / (ch.visit + 1))


def select_child_ucb1(node):
    """node.py:42-47 with mode 'normal': first maximum in child order."""
    ucb1_scores(node)
    best = node.children[0]
    for ch in node.children[1:]:
        if ch.ucb > best.ucb:
            best = ch
    return best


def rollout(g, board, node, rng=np.random):
    """ai/mcts.py:62-79 (MCTS.simulate): uniformly random legal moves to the end of the game on a copy of the
    board; +1 if the player who moved INTO `node` wins, -1 if the side to move at `node` wins, 0 for a draw."""
    sim = board.copy()
    player, mc, winner = node.player, node.move_count, -1
    while winner == -1 and mc < g.state_dim:
        moves = valid_moves(g, sim)
        a = moves[rng.randint(len(moves))]
        player = make_move(g, sim, player, a)
        mc += 1
        winner = check_winner(g, sim, 1 - player, a)
    if winner != -1:
        return 1 if winner == (1 - node.player) else -1
    return 0


def mcts_vanilla(g, board, root, iterations, simulate=None, rng=np.random):
    """ai/mcts.py:11-60 with model=None: UCB1 selection, children without priors, result = simulate(...)
    (default: the random rollout above; tests inject a deterministic function of the leaf position the way the
    golden generator replaces MCTS.simulate in the reference)."""
    for _ in range(iterations):
        node = root
        trace = [root]
        while node.children:
            node = select_child_ucb1(node)
            trace.append(node)
            make_move(g, board, 1 - node.player, node.action)
        if node.parent is not None:
            if check_winner(g, board, 1 - node.player, node.action) != -1:
                backup(g, trace, 1, board)
                continue
            if node.move_count == g.state_dim:
                backup(g, trace, 0, board)
                continue
        moves = valid_moves(g, board)
        for mv in moves:                                   # node.py:58: no prior
            node.children.append(Node(node, mv, 1 - node.player, node.move_count + 1))
        result = simulate(g, board, node) if simulate is not None else rollout(g, board, node, rng)
        backup(g, trace, result, board)


def hash_simulate(seed, qbits=16):
    """Deterministic stand-in for the rollout in parity runs: minus the hash evaluator's value of the leaf
    (the same number the GPU's hash evaluator backs up, ai/mcts.py:56 sign convention)."""
    def f(g, board, node):
        _, value = hash_eval_from_key(g, hash_eval_key(g, canonical_board(g, board, node.player), seed), qbits)
        return -float(value)
    return f


def expand(g, node, moves, priors):
    """node.py:50-59: one child per legal move in list order; prior gathered from the
    UNMASKED, un-renormalised policy vector by action index."""
    for mv in moves:
        node.children.append(Node(node, mv, 1 - node.player, node.move_count + 1,
                                  priors[action_idx(g, mv)]))


def backup(g, trace, value, board):
    """node.py:62-74: leaf->root, visit+=1, W+=v, v=-v, and undo the move on the shared board."""
    for nd in reversed(trace):
        nd.visit += 1
        nd.value += value
        value *= -1
        if nd.parent is not None:
            undo_move(g, board, nd.player, nd.action)


def mcts(g, evaluator, board, root, iterations, noise=None, stats=None):
    """ai/mcts.py:11-60, network mode.  `noise`: None = dirichlet off; else an f64[A]
    vector playing the role of np.random.dirichlet(...) at the root (utils.py:12-27:
    0.75*p + 0.25*noise, f64, not renormalised).  Terminal leaves are re-detected on
    every visit and never expanded; the root is never terminal-checked (mcts.py:25)."""
    for _ in range(iterations):
        node = root
        trace = [root]
        while node.children:
            node = select_child(node)
            trace.append(node)
            make_move(g, board, 1 - node.player, node.action)
        if node.parent is not None:
            if check_winner(g, board, 1 - node.player, node.action) != -1:
                backup(g, trace, 1, board)
                if stats is not None:
                    stats["terminal"] = stats.get("terminal", 0) + 1
                continue
            if node.move_count == g.state_dim:
                backup(g, trace, 0, board)
                if stats is not None:
                    stats["terminal"] = stats.get("terminal", 0) + 1
                continue
        moves = valid_moves(g, board)
        priors, value = evaluator(canonical_board(g, board, node.player))
        if node.parent is None and noise is not None:
            priors = (1 - 0.25) * priors + 0.25 * noise   # utils.py:24-25 -> f64
        expand(g, node, moves, priors)
        if stats is not None:
            stats["expand"] = stats.get("expand", 0) + 1
            stats["children"] = stats.get("children", 0) + len(moves)
            stats["depth"] = stats.get("depth", 0) + len(trace) - 1
        backup(g, trace, -float(value), board)           # mcts.py:56: result = -value.item()


def visit_policy(g, root):
    """utils.py:46-55: pi[a] = n_a / sum n, f64[A], zeros for moves without a child."""
    counts = np.zeros(g.action_dim)
    for ch in root.children:
        counts[action_idx(g, ch.action)] = ch.visit
    return counts / np.sum(counts)


def max_visit_child(root):
    """node.py:76-81: first maximum in child order."""
    best = root.children[0]
    for ch in root.children[1:]:
        if ch.visit > best.visit:
            best = ch
    return best


def sample_child(g, root, rng=np.random):
    """node.py:83-93: np.random.choice over action-indexed children with p = pi."""
    p = visit_policy(g, root)
    slots = [None] * g.action_dim
    for ch in root.children:
        slots[action_idx(g, ch.action)] = ch
    return rng.choice(slots, p=p)


def root_summary(g, root):
    """Flat view of a searched root used by golden vectors and parity tests."""
    return {
        "actions": [action_idx(g, ch.action) for ch in root.children],
        "moves": [list(map(int, ch.action)) for ch in root.children],
        "visits": [int(ch.visit) for ch in root.children],
        "values": [float(ch.value) for ch in root.children],
        "priors": [float(ch.prior) for ch in root.children],
        "root_visit": int(root.visit),
        "root_value": float(root.value),
    }


# --------------------------------------------------------------------------------------
# Self-play driver and trajectory -> buffer transform.
# --------------------------------------------------------------------------------------
def self_play(g, evaluator, iterations, noise_fn=None, rng=np.random, sample_plies=None,
              forced_moves=None, on_root=None):
    """gomoku.py:123-164 / connect4.py:117-151 / tictactoe.py:99-133 with a model.
    Fresh root every ply (no subtree reuse).  Move choice: Gomoku samples from pi while
    move_count < 8 then argmax; the 3-plane games always sample (sample_plies=None means
    'use the game's reference rule').  forced_moves teacher-forces the trajectory.
    Returns (boards, actions, policies, qs, winner) in the Gomoku 5-tuple layout."""
    board = new_board(g)
    player, mc = 0, 0
    boards, actions, policies, qs = [], [(-1, -1)], [], []
    if sample_plies is None:
        sample_plies = 8 if g.gid == 2 else g.state_dim + 1
    while True:
        root = Node(None, None, player, mc)
        mcts(g, evaluator, board, root, iterations,
             noise=None if noise_fn is None else noise_fn(g.action_dim))
        if on_root is not None:
            on_root(mc, board, root)
        policies.append(visit_policy(g, root))
        boards.append(board.copy())
        qs.append(root.value / root.visit)                # gomoku.py:141
        if forced_moves is not None:
            mv = tuple(forced_moves[mc])
            child = next(ch for ch in root.children if tuple(ch.action) == mv)
        elif mc < sample_plies:
            child = sample_child(g, root, rng)
        else:
            child = max_visit_child(root)
        player = make_move(g, board, player, child.action)
        mc += 1
        actions.append(child.action)
        winner = check_winner(g, board, root.player, child.action)
        if winner != -1:
            break
        if mc == g.state_dim:
            winner = -1
            break
    return boards, actions, policies, qs, winner


def rotate_pair(board, policy, k):
    """train.py:8-15."""
    _, rows, cols = board.shape
    return (np.rot90(board, k=k, axes=(1, 2)).copy(),
            np.rot90(policy.reshape(rows, cols), k=k).flatten())


def flip_pair(board, policy, mode):
    """train.py:17-27."""
    _, rows, cols = board.shape
    p2 = policy.reshape(rows, cols)
    if mode == "lr":
        return np.flip(board, axis=2).copy(), np.flip(p2, axis=1).flatten()
    return np.flip(board, axis=1).copy(), np.flip(p2, axis=0).flatten()


def trajectory_to_samples(g, boards, policies, winner, reward):
    """train.py:30-49: ply i belongs to side i%2; z = +reward for the winner's plies else
    -reward; canonical board; plies 0,1 once, later plies x8 in the order rot0, rot0-lr,
    rot0-tb, rot90, rot90-lr, rot90-tb, rot180, rot270.  Square boards with A == rows*cols
    only (the reference reshapes the policy to (rows, cols)); Connect4 gets the identity
    plus lr-flip as a documented build extension."""
    out = []
    side = 0
    for i in range(len(boards)):
        z = [reward] if side == winner else [-reward]
        cb = canonical_board(g, boards[i], side)
        side = 1 - side
        if i < 2:
            out.append((cb, policies[i], z))
            continue
        if g.gid == 1:
            out.append((cb, policies[i], z))
            out.append((np.flip(cb, axis=2).copy(), policies[i][::-1].copy(), z))
            continue
        for r in range(4):
            br, pr = rotate_pair(cb, policies[i], r)
            out.append((br, pr, z))
            if r < 2:
                out.append(flip_pair(br, pr, "lr") + (z,))
                out.append(flip_pair(br, pr, "tb") + (z,))
    return out
