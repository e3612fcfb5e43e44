"""Importable alias of the package directory `alpha-zero_b200/` (a hyphen is not a valid
module name).  Executes the real package's __init__ with its directory as __path__."""
import os as _os

_real = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))), "alpha-zero_b200")
__path__ = [_real]
with open(_os.path.join(_real, "__init__.py")) as _f:
    exec(compile(_f.read(), _os.path.join(_real, "__init__.py"), "exec"))
