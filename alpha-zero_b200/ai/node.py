"""Node view objects (reference: ai/node.py:20-97).

The search tree itself lives in HBM (csrc/az_forest.h); after MCTS.mcts returns, the root and
its children are materialised as these objects, in the reference's child order, with the
fields callers read: visit, value (W), prior, prevAction, currentPlayer, move_count, ucb.
"""
import math

import numpy as np

from .. import utils


class Node:
    def __init__(self, parent, prevAction, currentPlayer, move_count, prior=0.0):
        self.parent = parent
        self.visit = 0
        self.value = 0
        self.ucb = np.inf
        self.prior = prior
        self.move_count = move_count
        self.prevAction = prevAction
        self.currentPlayer = currentPlayer
        self.children = []

    def max_visit_child(self):
        """node.py:76-81: first maximum in child order."""
        return max(self.children, key=lambda ch: ch.visit)

    def sample_child(self, Game):
        """node.py:83-93: np.random.choice over action-indexed children with p = pi."""
        p = utils.get_probablity_distribution_of_children(self, Game)
        by_action = [None] * Game.action_dim
        for ch in self.children:
            by_action[Game.get_action_idx(ch.prevAction)] = ch
        return np.random.choice(by_action, p=p)

    def refresh_ucb(self, mode="network"):
        """PUCT ('network') or UCB1 ('normal') of the children as the reference would show them (utils.py:29-44)."""
        for ch in self.children:
            if mode == "normal":
                u = math.sqrt(2 * math.log(self.visit) / (ch.visit + 1)) if self.visit > 0 else 0.0
            else:
                u = ch.prior * math.sqrt(self.visit) / (ch.visit + 1)
            ch.ucb = u if ch.visit == 0 else (ch.value / ch.visit) + u

    def to_string(self, Game):
        p = utils.get_probablity_distribution_of_children(self.parent, Game)
        return (f"Node: {self.prevAction}, Value: {self.value}, Visit: {self.visit}, "
                f"P(Visit): {p[Game.get_action_idx(self.prevAction)]}, UCB: {self.ucb}")
