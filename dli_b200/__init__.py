"""Importable alias of the `distributed-llm-inferencing_b200/` package directory."""
import os as _os

_pkg = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))),
                     "distributed-llm-inferencing_b200")
__path__.insert(0, _pkg)
with open(_os.path.join(_pkg, "__init__.py")) as _f:
    exec(compile(_f.read(), _os.path.join(_pkg, "__init__.py"), "exec"))
del _f
