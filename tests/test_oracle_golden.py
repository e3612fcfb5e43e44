"""The CPU oracle (oracle/az_oracle.py) against outputs of the unmodified reference
(tests/golden/*, made by oracle/gen_golden.py).  CPU only."""
import numpy as np
import pytest

import az_oracle as orc
from golden_util import load_json, spec_of, board_from_words, position_after, fhex, GOLDEN


@pytest.mark.parametrize("name,g", [("rules_tictactoe.json", orc.TICTACTOE),
                                    ("rules_connect4.json", orc.CONNECT4),
                                    ("rules_gomoku7.json", orc.GOMOKU),
                                    ("rules_gomoku15.json", orc.gomoku_spec(15))])
def test_rules_playouts(name, g):
    n = 0
    for plies in load_json(name):
        board = orc.new_board(g)
        for rec in plies:
            ref = board_from_words(g, rec["p0"], rec["p1"])
            assert np.array_equal(ref[:2], board[:2])
            legal = orc.valid_moves(g, board)
            assert [r * g.cols + c for r, c in legal] == rec["legal"]
            mv = tuple(rec["move"])
            before = board.copy()
            nxt = orc.make_move(g, board, rec["player"], mv)
            assert nxt == 1 - rec["player"]
            assert orc.check_winner(g, board, rec["player"], mv) == rec["winner"]
            probe = board.copy()
            orc.undo_move(g, probe, nxt, mv)              # make/undo round trip
            assert np.array_equal(probe[:2], before[:2])
            n += 1
    assert n > 50


def test_gomoku_set_order_matches_real_set():
    rng = np.random.default_rng(5)
    for g in (orc.GOMOKU, orc.gomoku_spec(15), orc.gomoku_spec(9)):
        cells = g.rows * g.cols
        for r in range(g.rows):
            for c in range(g.cols):
                assert orc.cpython_tuple2_hash(r, c) == hash((r, c)) & orc.MASK64
        for _ in range(300):
            b = orc.new_board(g)
            k = int(rng.integers(0, cells))
            for i, cell in enumerate(rng.permutation(cells)[:k]):
                b[i % 2, cell // g.cols, cell % g.cols] = 1
            assert orc.valid_moves(g, b) == orc.valid_moves_pyset(g, b)


def test_gomoku_set_order_is_home_slot_order_when_homes_are_distinct():
    """What the device shortcut in csrc/az_rules.cuh relies on, checked on REAL CPython sets built with the reference's
    insertion sequence (gomoku.py:95-104): the final table has 8 / 32 / 128 slots for <= 4 / 5-18 / 19-76 keys, and whenever
    the keys' home slots hash & mask in that table are pairwise distinct, list(set) is the keys sorted by home slot."""
    rng = np.random.default_rng(11)
    hit = miss = 0
    for g in (orc.GOMOKU, orc.gomoku_spec(8), orc.gomoku_spec(6)):
        cells = g.rows * g.cols
        for _ in range(1500):
            b = orc.new_board(g)
            k = int(rng.integers(1, cells))
            for i, cell in enumerate(rng.permutation(cells)[:k]):
                b[i % 2, cell // g.cols, cell % g.cols] = 1
            real = set()
            for cell in orc.gomoku_insertion_sequence(g, b):
                real.add(divmod(cell, g.cols))
            if not real:
                continue
            n = len(real)
            mask = 7 if n <= 4 else (31 if n <= 18 else 127)
            homes = {a: hash(a) & mask for a in real}
            if len(set(homes.values())) < n:
                miss += 1
                continue
            hit += 1
            assert list(real) == sorted(real, key=lambda a: homes[a])
    assert hit > 1000 and miss > 50                           # both branches of the device code occur
    h7 = {hash((r, c)) & 127 for r in range(7) for c in range(7)}
    assert len(h7) == 49                                      # 7x7: every position with >= 19 candidates takes the shortcut


def test_puct_bits():
    for case in load_json("puct.json"):
        parent = orc.Node(None, None, 0, 0)
        parent.visit = case["parent_visit"]
        for j, ch in enumerate(case["children"]):
            p = fhex(ch["prior"])
            prior = np.float64(p) if case["prior_is_f64"] else np.float32(p)
            nd = orc.Node(parent, (0, j), 1, 1, prior)
            nd.visit = ch["visit"]
            nd.value = fhex(ch["value"]) if ch["visit"] else 0
            parent.children.append(nd)
        sel = orc.select_child(parent)
        assert parent.children.index(sel) == case["selected"]
        for nd, u, is32 in zip(parent.children, case["ucb"], case["ucb_is_f32"]):
            assert float(nd.ucb) == fhex(u)
            assert isinstance(nd.ucb, np.float32) == is32


def test_search_cases():
    cases = load_json("search.json")
    assert len(cases) > 100
    for case in cases:
        g = spec_of(case)
        board, player, mc = position_after(g, case["opening"])
        before = board.copy()
        root = orc.Node(None, None, player, mc)
        noise = None if case["noise"] is None else np.asarray(case["noise"], dtype=np.float64)
        orc.mcts(g, orc.HashEvaluator(g, case["seed"], case["qbits"]), board, root, case["iters"], noise=noise)
        assert np.array_equal(board, before)
        got, ref = orc.root_summary(g, root), case["result"]
        assert got["actions"] == ref["actions"]
        assert got["visits"] == ref["visits"]
        assert got["values"] == [fhex(v) for v in ref["values"]]
        assert got["priors"] == [fhex(v) for v in ref["priors"]]
        assert got["root_visit"] == ref["root_visit"] == case["iters"]
        assert got["root_value"] == fhex(ref["root_value"])


def test_vanilla_search_cases():
    """model=None (UCB1, children without priors; ai/mcts.py:57-59, utils.py:36,43): the reference's loop with
    MCTS.simulate replaced by a deterministic function of the leaf (oracle/gen_golden.py: ref_search_vanilla)."""
    cases = load_json("search_vanilla.json")
    assert len(cases) >= 30
    for case in cases:
        g = spec_of(case)
        board, player, mc = position_after(g, case["opening"])
        before = board.copy()
        root = orc.Node(None, None, player, mc)
        orc.mcts_vanilla(g, board, root, case["iters"], simulate=orc.hash_simulate(case["seed"], case["qbits"]))
        assert np.array_equal(board, before)
        ref = case["result"]
        assert [orc.action_idx(g, ch.action) for ch in root.children] == ref["actions"]
        assert [ch.visit for ch in root.children] == ref["visits"]
        assert [float(ch.value) for ch in root.children] == [fhex(v) for v in ref["values"]]
        assert root.visit == ref["root_visit"] == case["iters"]
        assert float(root.value) == fhex(ref["root_value"])


def test_rollout_statistics():
    """The random rollout (ai/mcts.py:62-79) from the empty TicTacToe board: first-player win / draw / loss
    frequencies of uniformly random play are known to be 58.5 % / 12.7 % / 28.8 %; 4000 rollouts must land
    within 3 points."""
    g = orc.TICTACTOE
    rng = np.random.RandomState(7)
    root = orc.Node(None, None, 0, 0)
    res = [orc.rollout(g, orc.new_board(g), root, rng) for _ in range(4000)]
    # root.player = 0 is to move, so +1 means player 1 (the one who "moved into" the root) wins
    lose, draw, win = res.count(1) / 4000, res.count(0) / 4000, res.count(-1) / 4000
    assert abs(win - 0.585) < 0.03 and abs(draw - 0.127) < 0.03 and abs(lose - 0.288) < 0.03, (win, draw, lose)


def test_selfplay_and_buffer_transform():
    recs = load_json("selfplay.json")
    arrs = np.load(f"{GOLDEN}/selfplay.npz")
    for i, rec in enumerate(recs):
        g = spec_of(rec)
        np.random.seed(rec["np_seed"])
        ev = orc.HashEvaluator(g, rec["seed"], rec["qbits"])
        boards, actions, policies, qs, winner = orc.self_play(
            g, ev, rec["iters"], noise_fn=lambda A: np.random.dirichlet([0.03] * A), rng=np.random)
        assert winner == rec["winner"] and len(boards) == rec["plies"]
        assert np.array_equal(np.stack(boards), arrs[f"t{i}_boards"])
        assert np.array_equal(np.stack(policies), arrs[f"t{i}_policies"])
        if f"t{i}_actions" in arrs:
            assert np.array_equal(np.asarray(actions, dtype=np.int64), arrs[f"t{i}_actions"])
            assert np.array_equal(np.asarray(qs), arrs[f"t{i}_qs"])
        if f"t{i}_buf_states" in arrs:
            reward = 0 if winner == -1 else 1
            samples = orc.trajectory_to_samples(g, boards, policies, winner, reward)
            assert np.array_equal(np.stack([s[0] for s in samples]), arrs[f"t{i}_buf_states"])
            assert np.array_equal(np.stack([s[1] for s in samples]), arrs[f"t{i}_buf_policies"])
            assert np.array_equal(np.asarray([s[2] for s in samples], dtype=np.float64), arrs[f"t{i}_buf_z"])


def test_fp32_sqrt_equals_rounded_f64_sqrt_for_every_visit_count():
    """The select kernel forms s = (float)sqrt((double)N_p) (utils.py:34,40 under NumPy 2) with one fp32
    square root.  Double rounding is innocuous for sqrt when 53 >= 2*24 + 2; checked for every N_p < 2^24."""
    n = np.arange(0, 1 << 24, dtype=np.int64)
    assert np.array_equal(np.sqrt(n.astype(np.float64)).astype(np.float32), np.sqrt(n.astype(np.float32)))
