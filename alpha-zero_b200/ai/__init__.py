from .nn import Net, B200Net  # noqa: F401
