"""Host driver of ONE self-play game through the reference-facing API (Game.mcts on a fresh
root per ply).  Follows games/gomoku.py:123-164 / connect4.py:117-151 / tictactoe.py:99-133.
The throughput path is train.collect_data, which plays many games concurrently on the GPU."""
import time

import numpy as np


def play_one_game(Game, board, model, mcts_iter, display, sample_until):
    """sample_until: plies that sample the move from the visit distribution (None = every ply,
    the 3-plane games' rule); later plies take the most visited child.  Without a model the
    most visited child is always taken."""
    from ..ai import Node
    from .. import utils
    player, ply = 0, 0
    boards, actions, policies, qs = [], [(-1, -1)], [], []
    t0 = time.time()
    while True:
        root = Node(None, None, player, ply)            # no subtree reuse
        Game.mcts(model, board, root, mcts_iter)
        policies.append(utils.get_probablity_distribution_of_children(root, Game))
        boards.append(board.copy())
        qs.append(root.value / root.visit)
        if model and (sample_until is None or ply < sample_until):
            child = root.sample_child(Game)
        else:
            child = root.max_visit_child()
        player = Game.make_move(board, player, child.prevAction)
        ply += 1
        actions.append(child.prevAction)
        winner = Game.check_winner(board, root.currentPlayer, child.prevAction)
        if winner != -1 or ply == Game.state_dim:
            if display:
                Game.display_board(board)
                if winner != -1:
                    print("time:", time.time() - t0, "s")
            return boards, actions, policies, qs, winner
