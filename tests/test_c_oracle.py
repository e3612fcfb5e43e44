"""The C oracle (oracle/az_oracle.c) against the reference's golden outputs and against the
Python oracle (incl. the build-defined virtual-loss mode).  CPU only."""
import numpy as np
import pytest

import az_oracle as orc
import c_oracle as co
from golden_util import load_json, spec_of, board_from_words, position_after, fhex


@pytest.mark.parametrize("name,g", [("rules_tictactoe.json", orc.TICTACTOE), ("rules_connect4.json", orc.CONNECT4),
                                    ("rules_gomoku7.json", orc.GOMOKU), ("rules_gomoku15.json", orc.gomoku_spec(15))])
def test_rules(name, g):
    for plies in load_json(name):
        for rec in plies:
            b = board_from_words(g, rec["p0"], rec["p1"])
            cells = co.cells_from_planes(b)
            assert co.legal(g, cells) == rec["legal"]
            r, c = rec["move"]
            cells[r * g.cols + c] = rec["player"] + 1
            assert co.winner(g, cells, rec["player"], r * g.cols + c) == rec["winner"]


def test_puct_bits():
    for case in load_json("puct.json"):
        pri = [fhex(c["prior"]) for c in case["children"]]
        vis = [c["visit"] for c in case["children"]]
        W = [fhex(c["value"]) for c in case["children"]]
        sel, sc = co.puct(case["parent_visit"], case["prior_is_f64"], pri, vis, W)
        assert sel == case["selected"]
        assert sc.tolist() == [fhex(u) for u in case["ucb"]]


def test_search_golden():
    for case in load_json("search.json"):
        g = spec_of(case)
        board, player, mc = position_after(g, case["opening"])
        noise = None if case["noise"] is None else np.asarray([case["noise"]], dtype=np.float64)
        r = co.search_batch(g, co.cells_from_planes(board)[None], [player], [mc], case["iters"], 1,
                            case["seed"], case["qbits"], noise)
        ref = case["result"]
        nc = int(r["n_children"][0])
        assert nc == len(ref["actions"])
        assert r["cells"][0, :nc].tolist() == [m[0] * g.cols + m[1] for m in ref["moves"]]
        assert r["visits"][0, :nc].tolist() == ref["visits"]
        assert r["W"][0, :nc].tolist() == [fhex(v) for v in ref["values"]]
        assert r["P"][0, :nc].tolist() == [fhex(v) for v in ref["priors"]]
        assert int(r["root_visit"][0]) == ref["root_visit"]
        assert float(r["root_value"][0]) == fhex(ref["root_value"])


def test_vanilla_search_golden():
    """UCB1 selection (model=None) against searches of the reference itself (tests/golden/search_vanilla.json)."""
    co.set_ucb1(True)
    try:
        for case in load_json("search_vanilla.json"):
            g = spec_of(case)
            board, player, mc = position_after(g, case["opening"])
            r = co.search_batch(g, co.cells_from_planes(board)[None], [player], [mc], case["iters"], 1, case["seed"], case["qbits"])
            ref = case["result"]
            nc = int(r["n_children"][0])
            assert nc == len(ref["actions"])
            assert r["visits"][0, :nc].tolist() == ref["visits"]
            assert r["W"][0, :nc].tolist() == [fhex(v) for v in ref["values"]]
            assert int(r["root_visit"][0]) == ref["root_visit"]
            assert float(r["root_value"][0]) == fhex(ref["root_value"])
    finally:
        co.set_ucb1(False)


def py_vl_search(g, ev, board, player, mc, iters, vl):
    """Python statement of the virtual-loss step machine the C oracle and the CUDA path
    implement (build-defined; no reference behaviour).  vl=1 is orc.mcts exactly."""
    root = orc.Node(None, None, player, mc)
    extra = {}                                             # id(node) -> [vloss, pending]

    def st(n):
        return extra.setdefault(id(n), [0, False])

    def score(ch, s32):
        v = st(ch)[0]
        n, W = ch.visit + v, ch.value - float(v)
        u = (ch.prior * s32) / np.float32(n + 1)
        return u if n == 0 else np.float32(W / n) + u

    sims = 0
    while sims < iters:
        pend, started, nterm = [], sims, 0
        while started < iters and len(pend) < vl and nterm < vl:       # a step closes after vl terminal descents too
            node, trace = root, [root]
            while node.children:
                s32 = np.float32(np.sqrt(np.float64(node.visit + st(node)[0])))
                best = node.children[0]
                bs = score(best, s32)
                for ch in node.children[1:]:
                    sc = score(ch, s32)
                    if sc > bs:
                        best, bs = ch, sc
                node = best
                trace.append(node)
                orc.make_move(g, board, 1 - node.player, node.action)
            if st(node)[1]:
                for nd in reversed(trace[1:]):
                    orc.undo_move(g, board, nd.player, nd.action)
                break
            term = None
            if node.parent is not None:
                if orc.check_winner(g, board, 1 - node.player, node.action) != -1:
                    term = 1
                elif node.move_count == g.state_dim:
                    term = 0
            if term is not None:
                orc.backup(g, trace, term, board)
                sims += 1
                started += 1
                nterm += 1
                continue
            pri, val = ev(orc.canonical_board(g, board, node.player))
            moves = orc.valid_moves(g, board)
            if vl > 1:
                st(node)[1] = True
                for nd in trace:
                    st(nd)[0] += 1
            for nd in reversed(trace[1:]):
                orc.undo_move(g, board, nd.player, nd.action)
            pend.append((trace, moves, pri, val))
            started += 1
        for trace, moves, pri, val in pend:
            leaf = trace[-1]
            orc.expand(g, leaf, moves, pri)
            st(leaf)[1] = False
            v = -float(val)
            for nd in reversed(trace):
                nd.visit += 1
                nd.value += v
                v = -v
                if vl > 1:
                    st(nd)[0] -= 1
            sims += 1
    return root


@pytest.mark.parametrize("g,vl,iters", [(orc.CONNECT4, 1, 150), (orc.CONNECT4, 4, 200), (orc.GOMOKU, 8, 300),
                                        (orc.TICTACTOE, 3, 120), (orc.GOMOKU, 1, 120)])
def test_vl_machine_matches_python(g, vl, iters):
    rng = np.random.default_rng(3)
    for trial in range(3):
        board = orc.new_board(g)
        player, mc = 0, 0
        for _ in range(int(rng.integers(0, 6))):
            mv = orc.valid_moves(g, board)
            player = orc.make_move(g, board, player, mv[int(rng.integers(len(mv)))])
            mc += 1
        seed = int(rng.integers(1 << 30))
        root = py_vl_search(g, orc.HashEvaluator(g, seed, 16), board.copy(), player, mc, iters, vl)
        r = co.search_batch(g, co.cells_from_planes(board)[None], [player], [mc], iters, vl, seed, 16)
        nc = int(r["n_children"][0])
        assert r["visits"][0, :nc].tolist() == [ch.visit for ch in root.children]
        assert r["W"][0, :nc].tolist() == [float(ch.value) for ch in root.children]
        assert int(r["root_visit"][0]) == root.visit == iters
