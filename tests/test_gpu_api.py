"""The reference-facing API on the GPU: MCTS.mcts / Game.self_play / save_data_to_buffer must
reproduce the unmodified reference's recorded outputs EXACTLY (same NumPy RNG stream, exact
priors through the shared hash evaluator)."""
import numpy as np
import pytest

import az_oracle as orc
from golden_util import load_json, spec_of, position_after, fhex, GOLDEN

pytestmark = pytest.mark.gpu


def game_cls(g):
    from alpha_zero_b200 import games
    G = {"tictactoe": games.TicTacToe, "connect4": games.Connect4, "gomoku": games.Gomoku}[g.name]
    if g.name == "gomoku":
        G.rows = G.cols = g.rows
        G.action_dim = G.state_dim = g.rows * g.cols
    return G


def test_mcts_call_matches_reference_search_cases(monkeypatch):
    from alpha_zero_b200.ai import MCTS, Node, HashEvaluator
    for case in load_json("search.json"):
        g = spec_of(case)
        G = game_cls(g)
        board, player, mc = position_after(g, case["opening"])
        before = board.copy()
        root = Node(None, None, player, mc)
        if case["noise"] is not None:
            monkeypatch.setattr(np.random, "dirichlet", lambda alpha, nz=case["noise"]: np.asarray(nz, dtype=np.float64))
        MCTS.mcts(HashEvaluator(case["seed"], case["qbits"]), board, root, G, case["iters"], dirichlet=case["noise"] is not None)
        monkeypatch.undo()
        ref = case["result"]
        assert np.array_equal(board, before)
        assert [list(ch.prevAction) for ch in root.children] == ref["moves"]
        assert [ch.visit for ch in root.children] == ref["visits"]
        assert [ch.value for ch in root.children] == [fhex(v) for v in ref["values"]]
        assert [float(ch.prior) for ch in root.children] == [fhex(v) for v in ref["priors"]]
        assert (root.visit, root.value) == (ref["root_visit"], fhex(ref["root_value"]))
        assert isinstance(root.children[0].prior, np.float64) == ref["prior_is_f64"]
    game_cls(orc.GOMOKU)


def test_self_play_and_buffer_match_reference_trajectories():
    from alpha_zero_b200.ai import HashEvaluator
    from alpha_zero_b200.replay_buffer import ReplayBuffer
    from alpha_zero_b200 import train
    recs = load_json("selfplay.json")
    arrs = np.load(f"{GOLDEN}/selfplay.npz")
    for i, rec in enumerate(recs):
        g = spec_of(rec)
        G = game_cls(g)
        np.random.seed(rec["np_seed"])
        out = G().self_play(HashEvaluator(rec["seed"], rec["qbits"]), rec["iters"])
        if len(out) == 5:
            boards, actions, policies, qs, winner = out
            assert np.array_equal(np.asarray(actions, dtype=np.int64), arrs[f"t{i}_actions"])
            assert np.array_equal(np.asarray(qs), arrs[f"t{i}_qs"])
        else:
            boards, policies, winner = out
            actions = qs = None
        assert winner == rec["winner"] and len(boards) == rec["plies"]
        assert np.array_equal(np.stack(boards), arrs[f"t{i}_boards"])
        assert np.array_equal(np.stack(policies), arrs[f"t{i}_policies"])
        if f"t{i}_buf_states" in arrs:
            buf = ReplayBuffer(100000)
            train.save_data_to_buffer(G, buf, (boards, actions, policies, qs, winner, 0 if winner == -1 else 1))
            assert np.array_equal(np.stack([e[0] for e in buf.buffer]), arrs[f"t{i}_buf_states"])
            assert np.array_equal(np.stack([e[1] for e in buf.buffer]), arrs[f"t{i}_buf_policies"])
            assert np.array_equal(np.asarray([e[2] for e in buf.buffer], dtype=np.float64), arrs[f"t{i}_buf_z"])
            s, p, z = buf.sample(min(16, buf.size()))
            assert s.dtype == p.dtype == z.dtype and str(s.dtype) == "torch.float32"
            assert tuple(p.shape) == (min(16, buf.size()), g.action_dim) and tuple(z.shape) == (min(16, buf.size()), 1)


def test_collect_data_and_train_step_run_end_to_end():
    import torch
    from alpha_zero_b200.ai import Net
    from alpha_zero_b200.games import Gomoku, Connect4
    from alpha_zero_b200.replay_buffer import ReplayBuffer
    from alpha_zero_b200 import train
    game_cls(orc.GOMOKU)
    torch.manual_seed(0)
    np.random.seed(0)
    for G, C in ((Gomoku, 2), (Connect4, 3)):
        net = Net((G.rows, G.cols), 3, 64, G.action_dim, 8, 1, C, dropout=0.1)
        buf = ReplayBuffer(100000)
        res = train.collect_data(G, net, buf, iterations=16, mcts_iter=32)
        assert sum(res) == 16 and buf.size() > 16 * 5
        s, p, z = buf.sample(32)
        assert np.allclose(p.sum(1).numpy(), 1.0, atol=1e-5) and set(np.unique(z.numpy())) <= {-1.0, 0.0, 1.0}
        before = [q.detach().clone() for q in net.parameters()]
        loss = train.train(net, 32, buf, 2, 2.5e-4, "cuda")
        assert np.isfinite(loss[0]) and any(not torch.equal(a, b.cpu()) for a, b in zip(before, net.parameters()))
        net.to("cpu")
        res2 = train.collect_data(G, net, buf, iterations=4, mcts_iter=16)      # weights changed: re-upload path
        assert sum(res2) == 4


def test_play_games_batched_with_refill_is_self_consistent():
    """train.play_games keeps `concurrent` games in flight and refills finished slots.  Every recorded
    game must replay under the scalar rules of the Game class: boards follow from the actions, the winner is
    what check_winner says, policies are visit counts / (iters - 1), moves are legal."""
    import torch
    from alpha_zero_b200.ai import Net
    from alpha_zero_b200.games import TicTacToe, Connect4, Gomoku
    from alpha_zero_b200 import train
    game_cls(orc.GOMOKU)
    torch.manual_seed(1)
    np.random.seed(1)
    for G, iters, n, conc in ((TicTacToe, 24, 40, 12), (Connect4, 24, 24, 24), (Gomoku, 16, 20, 8)):
        net = Net((G.rows, G.cols), 3, 64, G.action_dim, 8, 1, G.feature_dim, dropout=0.1)
        games, visits = train.play_games(G, net, n, iters, dirichlet=True, return_visits=True, concurrent=conc)
        assert len(games) == n
        for (boards, actions, policies, qs, winner), vis in zip(games, visits):
            assert len(boards) == len(policies) == len(qs) == len(actions) - 1 == len(vis) >= 1
            b = G().board.copy()
            player, w = 0, -1
            for i, mv in enumerate(actions[1:]):
                assert np.array_equal(boards[i], b)
                assert tuple(mv) in [tuple(m) for m in G.get_valid_moves(b)]
                assert vis[i].sum() == iters - 1 and np.allclose(policies[i], vis[i] / vis[i].sum())
                assert vis[i][G.get_action_idx(mv)] > 0
                mover = player
                player = G.make_move(b, mover, mv)
                w = G.check_winner(b, mover, mv)
                assert (w != -1 or i + 1 == G.state_dim) == (i == len(boards) - 1)
            assert winner == w
