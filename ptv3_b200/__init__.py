"""Importable alias of the product package.

The product lives in `pointcept-keypointdetection_b200/` (the directory name the
build contract fixes; not a valid Python identifier), so this package points
its search path there and executes that directory's `__init__.py`.
"""
import os as _os

_real = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))),
                      "pointcept-keypointdetection_b200")
__path__ = [_real]
with open(_os.path.join(_real, "__init__.py")) as _f:
    exec(compile(_f.read(), _os.path.join(_real, "__init__.py"), "exec"))
del _f
