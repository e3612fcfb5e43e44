#!/usr/bin/env python
"""Generate tests/golden/* by EXECUTING THE UNMODIFIED REFERENCE (build container only).

The reference (/root/reference, pure Python) ships no tests or golden vectors, so the pins
for the oracle and for the CUDA path are outputs of the reference itself.  Importing the
reference's `games` package writes a logs/ directory next to it, hence we run a scratch
copy under a temp dir.  Nothing here runs on the GPU box; the fixtures are committed.

Shims applied OUTSIDE the reference copy (reference gaps):
  * TicTacToe/Connect4 get `get_canonical_board` = swap planes 0/1, keep plane 2 -- the
    author's own legacy code in the comments of ai/mcts.py:126-137;
  * the evaluator is injected through the reference's `MCTS.cache` hook
    (ai/mcts.py:40-41): a dict subclass that computes the hash evaluator from the key
    bytes, so the reference's loop runs unmodified and never calls a model;
  * `np.random.dirichlet` is patched to return a recorded vector where noise is on.

Usage: python oracle/gen_golden.py [--ref /root/reference] [--out tests/golden]
"""
import argparse
import json
import os
import shutil
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import az_oracle as orc  # noqa: E402


def load_reference(ref_src):
    tmp = tempfile.mkdtemp(prefix="az_ref_")
    dst = os.path.join(tmp, "ref")
    shutil.copytree(ref_src, dst)
    sys.path.insert(0, dst)          # must be first: reference's test.py shadows stdlib `test`
    import games as rgames           # noqa
    import ai as rai                 # noqa
    import utils as rutils           # noqa
    import train as rtrain           # noqa
    import replay_buffer as rbuf     # noqa
    from ai.nn import Net as RNet    # noqa

    def legacy_canonical(board, current_player):      # ai/mcts.py:126-137 (commented legacy)
        if current_player == 0:
            return board
        out = np.empty_like(board)
        out[0], out[1], out[2] = board[1], board[0], board[2]
        return out

    for cls in (rgames.TicTacToe, rgames.Connect4):
        cls.get_canonical_board = staticmethod(legacy_canonical)
        cls.feature_dim = 3
    return dict(games=rgames, ai=rai, utils=rutils, train=rtrain, buf=rbuf, Net=RNet, tmp=tmp)


class HashEvalCache(dict):
    """Drop-in for MCTS.cache: every key 'hits' and the value is computed from the key."""

    def __init__(self, g, seed, qbits):
        super().__init__()
        self.g, self.seed, self.qbits = g, seed, qbits

    def __contains__(self, key):
        return True

    def __getitem__(self, key):
        canon = np.frombuffer(key, dtype=np.float32).reshape(self.g.planes, self.g.rows, self.g.cols)
        pri, val = orc.hash_eval_from_key(self.g, orc.hash_eval_key(self.g, canon, self.seed), self.qbits)
        return pri, val               # np.float32 scalar has .item() like the tensor it replaces


def ref_game_class(R, g):
    G = {0: R["games"].TicTacToe, 1: R["games"].Connect4, 2: R["games"].Gomoku}[g.gid]
    if g.gid == 2:
        G.rows, G.cols = g.rows, g.cols
        G.action_dim = G.state_dim = g.rows * g.cols
    return G


def words_hex(g, plane):
    return ["%016x" % w for w in orc.board_words(g, plane)]


def gen_rules(R, g, n_playouts, rng):
    G = ref_game_class(R, g)
    playouts = []
    for _ in range(n_playouts):
        board = np.zeros((g.planes, g.rows, g.cols), dtype=np.float32)
        player, mc, plies = 0, 0, []
        while True:
            legal = G.get_valid_moves(board)
            mv = legal[int(rng.integers(len(legal)))]
            rec = {"p0": words_hex(g, board[0]), "p1": words_hex(g, board[1]),
                   "legal": [int(r) * g.cols + int(c) for r, c in legal],
                   "move": [int(mv[0]), int(mv[1])], "player": player}
            nxt = G.make_move(board, player, mv)
            mc += 1
            w = G.check_winner(board, player, mv)
            rec["winner"] = int(w)
            plies.append(rec)
            player = nxt
            if w != -1 or mc == G.state_dim:
                break
        playouts.append(plies)
    return playouts


def setup_position(R, g, opening):
    G = ref_game_class(R, g)
    board = np.zeros((g.planes, g.rows, g.cols), dtype=np.float32)
    player = 0
    for mv in opening:
        player = G.make_move(board, player, tuple(mv))
    return G, board, player, len(opening)


def random_opening(R, g, n, rng):
    """n random legal plies without finishing the game (retry on early wins)."""
    G = ref_game_class(R, g)
    while True:
        board = np.zeros((g.planes, g.rows, g.cols), dtype=np.float32)
        player, moves, ok = 0, [], True
        for _ in range(n):
            legal = G.get_valid_moves(board)
            mv = legal[int(rng.integers(len(legal)))]
            G.make_move(board, player, mv)
            if G.check_winner(board, player, mv) != -1:
                ok = False
                break
            moves.append([int(mv[0]), int(mv[1])])
            player = 1 - player
        if ok:
            return moves


def ref_search(R, g, opening, iters, seed, qbits, noise):
    G, board, player, mc = setup_position(R, g, opening)
    MCTS, Node = R["ai"].MCTS, R["ai"].Node
    MCTS.cache = HashEvalCache(g, seed, qbits)
    before = board.copy()
    root = Node(None, None, player, mc)
    saved = np.random.dirichlet
    if noise is not None:
        np.random.dirichlet = lambda alpha: np.asarray(noise, dtype=np.float64)
    try:
        MCTS.mcts(True, board, root, G, iters, dirichlet=noise is not None)
    finally:
        np.random.dirichlet = saved
    assert np.array_equal(before, board), "reference must restore the board"
    return {
        "actions": [int(G.get_action_idx(ch.prevAction)) for ch in root.children],
        "moves": [[int(ch.prevAction[0]), int(ch.prevAction[1])] for ch in root.children],
        "visits": [int(ch.visit) for ch in root.children],
        "values": [float(ch.value).hex() for ch in root.children],
        "priors": [float(ch.prior).hex() for ch in root.children],
        "prior_is_f64": bool(root.children and isinstance(root.children[0].prior, np.float64)),
        "root_visit": int(root.visit),
        "root_value": float(root.value).hex(),
    }


def ref_search_vanilla(R, g, opening, iters, seed, qbits=16):
    """The reference's vanilla mode (model=None: UCB1 + expansion without priors, ai/mcts.py:57-59, utils.py:36,43)
    with its loop unmodified; only the RNG-bound rollout MCTS.simulate is replaced by a deterministic function of
    the leaf position (minus the hash evaluator's value), so the search is reproducible."""
    G, board, player, mc = setup_position(R, g, opening)
    MCTS, Node = R["ai"].MCTS, R["ai"].Node
    saved = MCTS.simulate

    def simulate(Game, b, node):
        canon = orc.canonical_board(g, b, node.currentPlayer)
        _, value = orc.hash_eval_from_key(g, orc.hash_eval_key(g, canon, seed), qbits)
        return -float(value)
    MCTS.simulate = staticmethod(simulate)
    before = board.copy()
    root = Node(None, None, player, mc)
    try:
        MCTS.mcts(None, board, root, G, iters)
    finally:
        MCTS.simulate = saved
    assert np.array_equal(before, board), "reference must restore the board"
    return {
        "actions": [int(G.get_action_idx(ch.prevAction)) for ch in root.children],
        "visits": [int(ch.visit) for ch in root.children],
        "values": [float(ch.value).hex() for ch in root.children],
        "root_visit": int(root.visit),
        "root_value": float(root.value).hex(),
    }


def gen_vanilla(R, rng):
    cases = []
    plan = [(orc.TICTACTOE, [0, 2, 5], [1, 2, 50, 400]), (orc.CONNECT4, [0, 4, 15, 30], [8, 200, 800]),
            (orc.GOMOKU, [0, 3, 12, 30], [2, 200, 800])]
    for g, opens, iters_list in plan:
        for n_open in opens:
            opening = random_opening(R, g, n_open, rng)
            for iters in iters_list:
                seed = int(rng.integers(1 << 40))
                cases.append({"game": g.name, "rows": g.rows, "opening": opening, "iters": iters, "seed": seed,
                              "qbits": 16, "result": ref_search_vanilla(R, g, opening, iters, seed)})
    return cases


def gen_puct(R, rng, n_cases):
    """utils.calcUcbOfChildrenFromParent('network') on synthetic nodes -> exact score bits."""
    Node, utils = R["ai"].Node, R["utils"]
    cases = []
    for _ in range(n_cases):
        nch = int(rng.integers(1, 12))
        f64 = bool(rng.integers(0, 4) == 0)
        parent = Node(None, None, 0, 0)
        visits = rng.integers(0, 40, size=nch)
        parent.visit = int(visits.sum()) + 1
        chs = []
        for j in range(nch):
            p32 = np.float32(rng.random() * 0.3)
            prior = np.float64(0.75) * p32 + 0.25 * rng.random() if f64 else p32
            ch = Node(parent, (0, j), 1, 1, prior)
            ch.visit = int(visits[j])
            w = 0.0
            for _k in range(ch.visit):                   # f64 sum of f32-valued terms
                w += float(np.float32(rng.uniform(-1, 1)))
            ch.value = w if ch.visit else 0
            parent.children.append(ch)
            chs.append({"prior": float(prior).hex(), "visit": ch.visit, "value": float(ch.value).hex()})
        utils.calcUcbOfChildrenFromParent(parent, "network")
        sel = parent.select("network")
        cases.append({
            "parent_visit": parent.visit, "prior_is_f64": f64, "children": chs,
            "ucb": [float(ch.ucb).hex() for ch in parent.children],
            "ucb_is_f32": [bool(isinstance(ch.ucb, np.float32)) for ch in parent.children],
            "selected": parent.children.index(sel),
        })
    return cases


def gen_selfplay(R, g, iters, seed, np_seed, qbits=16):
    """Full reference self_play (Dirichlet ON, np.random sampling) + save_data_to_buffer."""
    G = ref_game_class(R, g)
    MCTS = R["ai"].MCTS
    MCTS.cache = HashEvalCache(g, seed, qbits)
    np.random.seed(np_seed)
    game = G()
    out = game.self_play(True, iters)
    if len(out) == 5:
        boards, actions, policies, qs, winner = out
    else:                                                # T3/C4: 3-tuple (connect4.py:151)
        boards, policies, winner = out
        actions, qs = None, None
    reward = 0 if winner == -1 else 1
    rec = {"game": g.name, "rows": g.rows, "iters": iters, "seed": seed, "np_seed": np_seed,
           "qbits": qbits, "winner": int(winner), "plies": len(boards)}
    arrays = {"boards": np.stack(boards), "policies": np.stack(policies)}
    if actions is not None:
        arrays["actions"] = np.asarray(actions, dtype=np.int64)
        arrays["qs"] = np.asarray(qs, dtype=np.float64)
    if g.gid == 2 or g.gid == 0:                         # square + A == rows*cols only
        buf = R["buf"].ReplayBuffer(100000)
        R["train"].save_data_to_buffer(G, buf, (boards, actions, policies, qs, winner, reward))
        arrays["buf_states"] = np.stack([e[0] for e in buf.buffer])
        arrays["buf_policies"] = np.stack([e[1] for e in buf.buffer])
        arrays["buf_z"] = np.asarray([e[2] for e in buf.buffer], dtype=np.float64)
    return rec, arrays


def deterministic_state_dict(model, seed):
    """Weights as a pure function of (parameter name, shape, seed) -- regenerated in tests."""
    import torch
    sd = {}
    for i, (k, v) in enumerate(sorted(model.state_dict().items())):
        rng = np.random.default_rng([seed, i])
        x = rng.standard_normal(tuple(v.shape)).astype(np.float32)
        if k.endswith("bias"):
            x *= 0.1
        elif "norm" in k and k.endswith("weight"):
            x = 1.0 + 0.1 * x
        elif k.endswith("cls_token") or k.endswith("pos_embedding"):
            x *= 0.5
        else:
            fan_in = int(np.prod(v.shape[1:])) if v.ndim > 1 else int(v.shape[0])
            x *= 1.0 / np.sqrt(fan_in)
        sd[k] = torch.from_numpy(x)
    return sd


def gen_net(R, cfgs, rng):
    import torch
    out_rec, out_arr = [], {}
    for ci, c in enumerate(cfgs):
        net = R["Net"](c["img"], c["patch"], c["dim"], c["A"], c["heads"], c["depth"], c["C"], dropout=0.1)
        net.load_state_dict(deterministic_state_dict(net, c["wseed"]))
        net.eval()
        x = (rng.random((c["B"], c["C"], c["img"], c["img"])) < 0.3).astype(np.float32)
        x[:, 1] *= (1 - x[:, 0])                         # no cell owned by both players
        if c["C"] == 3:                                  # plane 2 is the constant side-to-move plane
            x[:, 2] = rng.integers(0, 2, c["B"]).astype(np.float32)[:, None, None]
        with torch.no_grad():
            logits, values = net(torch.from_numpy(x))
        out_rec.append(dict(c, keys=sorted(net.state_dict().keys()),
                            shapes=[list(net.state_dict()[k].shape) for k in sorted(net.state_dict().keys())]))
        out_arr[f"x{ci}"] = x
        out_arr[f"logits{ci}"] = logits.numpy()
        out_arr[f"values{ci}"] = values.numpy()
    return out_rec, out_arr


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ref", default="/root/reference")
    ap.add_argument("--out", default=os.path.join(os.path.dirname(HERE), "tests", "golden"))
    ap.add_argument("--only", default="", help="'vanilla': write only search_vanilla.json (own RNG stream)")
    args = ap.parse_args()
    os.makedirs(args.out, exist_ok=True)
    R = load_reference(args.ref)
    rng = np.random.default_rng(20261104)

    def dump(name, obj):
        with open(os.path.join(args.out, name), "w") as f:
            json.dump(obj, f, separators=(",", ":"))
        print("wrote", name, os.path.getsize(os.path.join(args.out, name)), "bytes")

    if args.only == "vanilla":
        dump("search_vanilla.json", gen_vanilla(R, np.random.default_rng(20261105)))
        return
    g15 = orc.gomoku_spec(15)
    # 1. rules ------------------------------------------------------------------------
    dump("rules_tictactoe.json", gen_rules(R, orc.TICTACTOE, 40, rng))
    dump("rules_connect4.json", gen_rules(R, orc.CONNECT4, 40, rng))
    dump("rules_gomoku7.json", gen_rules(R, orc.GOMOKU, 40, rng))
    dump("rules_gomoku15.json", gen_rules(R, g15, 4, rng))
    # 2. search -----------------------------------------------------------------------
    cases = []
    plan = [(orc.TICTACTOE, [0, 1, 2, 4, 6], [1, 2, 30, 100, 400]),
            (orc.CONNECT4, [0, 1, 3, 6, 12, 20, 30], [1, 8, 100, 800]),
            (orc.GOMOKU, [0, 1, 2, 5, 9, 16, 30], [1, 2, 100, 800]),
            (g15, [0, 3, 20], [200])]
    for g, opens, iters_list in plan:
        for n_open in opens:
            opening = random_opening(R, g, n_open, rng)
            for iters in iters_list:
                for variant in range(2):
                    seed = int(rng.integers(1 << 40))
                    qbits = 16 if variant == 0 else 3     # 3 bits => many exact prior ties
                    noise = None
                    if variant == 1 and iters >= 100:
                        noise = rng.dirichlet([0.03] * g.action_dim).tolist()
                    res = ref_search(R, g, opening, iters, seed, qbits, noise)
                    cases.append({"game": g.name, "rows": g.rows, "opening": opening, "iters": iters,
                                  "seed": seed, "qbits": qbits, "noise": noise, "result": res})
    dump("search.json", cases)
    dump("search_vanilla.json", gen_vanilla(R, np.random.default_rng(20261105)))
    # 3. PUCT arithmetic ----------------------------------------------------------------
    dump("puct.json", gen_puct(R, rng, 400))
    # 4. self-play trajectories + buffer transform ---------------------------------------
    recs, arrs = [], {}
    for i, (g, iters) in enumerate([(orc.GOMOKU, 48), (orc.GOMOKU, 120), (orc.TICTACTOE, 60), (orc.CONNECT4, 64)]):
        rec, arr = gen_selfplay(R, g, iters, seed=1000 + i, np_seed=7 + i)
        recs.append(rec)
        for k, v in arr.items():
            arrs[f"t{i}_{k}"] = v
    dump("selfplay.json", recs)
    np.savez_compressed(os.path.join(args.out, "selfplay.npz"), **arrs)
    # 5. Net forward ----------------------------------------------------------------------
    cfgs = [dict(img=7, patch=3, dim=64, A=49, heads=8, depth=1, C=2, B=6, wseed=11),
            dict(img=7, patch=5, dim=128, A=49, heads=8, depth=2, C=2, B=5, wseed=12),
            dict(img=3, patch=3, dim=64, A=9, heads=4, depth=2, C=3, B=4, wseed=13),
            dict(img=7, patch=5, dim=512, A=49, heads=8, depth=1, C=2, B=3, wseed=14)]
    nrec, narr = gen_net(R, cfgs, rng)
    dump("net.json", nrec)
    np.savez_compressed(os.path.join(args.out, "net.npz"), **narr)
    shutil.rmtree(R["tmp"], ignore_errors=True)


if __name__ == "__main__":
    main()
