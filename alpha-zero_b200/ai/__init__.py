from .nn import Net, B200Net  # noqa: F401
from .node import Node  # noqa: F401
from .mcts import MCTS, HashEvaluator  # noqa: F401
