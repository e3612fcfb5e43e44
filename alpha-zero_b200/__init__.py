"""alpha-zero_b200: B200-native self-play hot path behind the reference's Game/MCTS/Net API.

Everything that computes lives in libaz_b200.so (hand-written sm_100a CUDA behind the C ABI
of include/az_b200.h).  The directory name carries a hyphen (as the task names it); import
it as `alpha_zero_b200` (the shim package next to it).
"""
from . import lib  # noqa: F401
from .forest import Forest, GameSpec, GAMES, game_spec  # noqa: F401

from .selfplay import SelfPlayEngine  # noqa: F401

__all__ = ["lib", "Forest", "GameSpec", "GAMES", "game_spec", "SelfPlayEngine"]
