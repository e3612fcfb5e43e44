"""Arena / gating with the reference's entry points (test.py:60-140), all games in flight at once.

    compete(Game, model1, model2, model1_mcts_iter, model2_mcts_iter, sampling, display) -> (winner, final_board)
    compare(Game, best_model, contender_model, best_iter, contender_iter, iterations, sampling, early_stopping)

The reference plays the `iterations` games of compare one after another, one batch-1 search per ply.  Games are
independent, so here every game is a tree of two GPU forests (one per model, each with its own evaluation cache and
step graph): at each ply the games whose mover is model A are searched on forest A and the others on forest B -- two
batched searches per ply for the whole arena.  Per game nothing changes: a fresh root per ply, Dirichlet noise ON
(test.py:82,84 call Game.mcts without `dirichlet`, whose default is True), the move = root.max_visit_child() (first
maximum in CHILD order, ai/node.py:76-81) or root.sample_child() while move_count < 20 when sampling (test.py:87),
sides swapped from game iterations//2 on (test.py:115-118), draws count half a win for both (test.py:124-126), and the
early-stopping rule of test.py:131-138 is applied to the per-game results in game order, which returns what the
sequential loop returns for the same outcomes.  Noise and sampling uniforms come from the global NumPy RNG.
"""
import numpy as np

from .ai.mcts import MCTS


def _play(Game, models, iters, first_is_a, sampling, rng=np.random, log=None):
    """n games; game i has model A moving first iff first_is_a[i].  Returns winners (0 / 1 / -1 as check_winner) and
    final boards.  models = (A, B), iters = (iters_a, iters_b)."""
    from . import train as _tr
    n = len(first_is_a)
    C, R, W, A = Game.feature_dim, Game.rows, Game.cols, Game.action_dim
    boards = np.zeros((n, C, R, W), np.float32)
    players = np.zeros(n, np.int32)
    plies = np.zeros(n, np.int32)
    alive = np.ones(n, bool)
    winners = np.full(n, -1, np.int64)
    first_is_a = np.asarray(first_is_a, bool)
    ply = 0
    while alive.any():
        a_moves = first_is_a == (ply % 2 == 0)                       # model A is player 0 in its "first" games
        visits = np.zeros((n, A), np.int64)
        best = np.full(n, -1, np.int64)
        noise_used = np.zeros((n, A))
        for which, mask in ((0, alive & a_moves), (1, alive & ~a_moves)):
            if not mask.any():
                continue
            model = models[which]
            noise = None
            if model:                                                # vanilla mode (model=None) never mixes noise in
                noise = np.zeros((n, A))
                for i in np.nonzero(mask)[0]:                        # one draw per searched root (utils.py:21), in game order
                    noise[i] = rng.dirichlet([0.03] * A)
                noise_used[mask] = noise[mask]
            v, _q, b = MCTS.mcts_batch(model, boards, players, plies, Game, iters[which], noise=noise, active=mask.astype(np.uint8),
                                       forest_tag=1 + which, return_best=True)
            visits[mask], best[mask] = v[mask], b[mask]
        live = np.nonzero(alive)[0]
        cells = best[live].copy()
        us = np.full(len(live), -1.0)
        if sampling and ply < 20:                                     # test.py:87: sample_child while move_count < 20
            for j, i in enumerate(live):
                p = visits[i] / np.sum(visits[i].astype(np.float64))
                us[j] = rng.random_sample()
                cdf = p.cumsum()
                cdf /= cdf[-1]
                a = int(cdf.searchsorted(us[j], side="right"))       # np.random.choice(children by action, p=p), node.py:92
                if a < A and visits[i, a] > 0:
                    if Game.name == "connect4":
                        col_free = (boards[i, 0, :, a] + boards[i, 1, :, a]) == 0
                        cells[j] = (int(col_free.sum()) - 1) * W + a
                    else:
                        cells[j] = a
        r, c = cells // W, cells % W
        movers = players[live].astype(np.int64)
        if log is not None:
            log.append(dict(ply=ply, live=live.copy(), noise=noise_used[live].copy(), cells=cells.copy(), u=us,
                            visits=visits[live].copy(), boards=boards[live].copy()))
        boards[live, movers, r, c] = 1
        if C == 3:
            boards[live, 2] = (1 - movers)[:, None, None]
        players[live] = 1 - movers
        plies[live] += 1
        won = _tr._wins(Game, boards[live], movers, r, c)
        over = won | (plies[live] == Game.state_dim)
        done = live[over]
        winners[done] = np.where(won[over], movers[over], -1)
        alive[done] = False
        ply += 1
    return winners, boards


def compete(Game, model1, model2, model1_mcts_iter=50, model2_mcts_iter=50, sampling=False, display=False):
    """test.py:60-104: one game, model1 moves first.  Returns (winner, final_board), winner = 0 / 1 / -1 (draw)."""
    winners, boards = _play(Game, (model1, model2), (model1_mcts_iter, model2_mcts_iter), [True], sampling)
    if display:
        Game.display_board(boards[0])
        print("It's a draw!" if winners[0] == -1 else f"Player {int(winners[0])} wins!")
    return int(winners[0]), boards[0]


def compare(Game, best_model, contender_model, best_model_mcts_iter, contender_model_mcts_iter, iterations, sampling, early_stopping,
            rng=np.random, log=None):
    """test.py:106-140.  Returns the contender's score rate, or 1 / 0 when early stopping decides (accept / reject)."""
    for m in (best_model, contender_model):
        if m is not None and hasattr(m, "eval"):
            m.eval()
    first_is_best = np.arange(iterations) < iterations // 2          # the best model moves first in the first half
    winners, boards = _play(Game, (best_model, contender_model), (best_model_mcts_iter, contender_model_mcts_iter), first_is_best,
                            sampling, rng=rng, log=log)
    win_count = [0, 0, 0]                                             # best, contender, draws
    for i in range(iterations):
        if winners[i] == 0:
            win_count[0 if first_is_best[i] else 1] += 1
        elif winners[i] == 1:
            win_count[1 if first_is_best[i] else 0] += 1
        else:
            win_count[0] += 0.5
            win_count[1] += 0.5
            win_count[2] += 1
        Game.logger.debug(f"compare iter({i + 1}/{iterations}) win count:{win_count}, best model win rate:{win_count[0] / (i + 1)}, "
                          f"contender model win rate:{win_count[1] / (i + 1)}, draw rate:{win_count[2] / (i + 1)}")
        if early_stopping:
            if win_count[1] >= int(iterations * 0.55):
                Game.logger.debug("compare early stopped, model accepted")
                return 1
            if win_count[1] + (iterations - (i + 1)) < int(iterations * 0.55):
                Game.logger.debug("compare early stopped, model rejected")
                return 0
    print("win count:", win_count, "best model win rate:", win_count[0] / iterations, "contender model win rate:",
          win_count[1] / iterations, "draw rate:", win_count[2] / iterations)
    return win_count[1] / iterations
