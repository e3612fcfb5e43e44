"""Helpers with the reference's names (utils.py): visit-count policy, Dirichlet mixing,
model checkpoints, per-game loggers."""
import logging
import os

import numpy as np


def add_dirichlet_noise(priors, epsilon=0.25, alpha=0.03):
    """utils.py:12-27: (1-eps)*p + eps*Dir(alpha), not renormalised.  The search kernels do
    the mixing on the device; this host version documents/draws the same noise."""
    noise = np.random.dirichlet([alpha] * len(priors))
    return (1 - epsilon) * priors + epsilon * noise


def get_probablity_distribution_of_children(node, Game):
    """utils.py:46-55: pi[a] = n_a / sum(n), float64[A], zero where no child exists."""
    counts = np.zeros(Game.action_dim)
    for ch in node.children:
        counts[Game.get_action_idx(ch.prevAction)] = ch.visit
    return counts / np.sum(counts)


def save_model(model, version, directory=None):
    """utils.py:57-61: state_dict -> models/model_v{N}.pth."""
    import torch
    directory = directory or os.path.join(os.getcwd(), "models")
    os.makedirs(directory, exist_ok=True)
    torch.save(model.state_dict(), os.path.join(directory, f"model_v{version}.pth"))


def load_model(model, name, directory=None):
    """utils.py:63-69."""
    import torch
    directory = directory or os.path.join(os.getcwd(), "models")
    model.load_state_dict(torch.load(os.path.join(directory, name)))
    return model


def get_game_logger(game_name):
    """utils.py:71-90 opens logs/<game>.log at import time; here a logger without side effects
    (attach handlers yourself, or set AZ_LOG_DIR to get the reference's file log)."""
    logger = logging.getLogger(game_name)
    log_dir = os.environ.get("AZ_LOG_DIR")
    if log_dir and not logger.handlers:
        os.makedirs(log_dir, exist_ok=True)
        h = logging.FileHandler(os.path.join(log_dir, f"{game_name}.log"), mode="a")
        h.setFormatter(logging.Formatter("%(asctime)s - %(name)s - %(levelname)s - %(message)s"))
        logger.addHandler(h)
        logger.setLevel(logging.DEBUG)
    elif not logger.handlers:
        logger.addHandler(logging.NullHandler())
    return logger
