"""Importable alias of the product package directory `gravo-mg_b200/` (a hyphen is not a valid
Python identifier, so `import gravomg_b200` resolves its submodules from that directory)."""
import os as _os

_PKG_DIR = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))), "gravo-mg_b200")
__path__ = [_PKG_DIR]
with open(_os.path.join(_PKG_DIR, "__init__.py")) as _f:
    exec(compile(_f.read(), _os.path.join(_PKG_DIR, "__init__.py"), "exec"))
