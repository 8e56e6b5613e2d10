"""Importable alias of the package directory `ray-project-ray_b200/`.

The product lives in `ray-project-ray_b200/` (the name the layout contract asks for); a
hyphen cannot appear in a Python module name, so this stub points `__path__` at that
directory and runs its `__init__.py`. `import ray_project_ray_b200` is the only thing it does.
"""
import os as _os

_real = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))), "ray-project-ray_b200")
__path__ = [_real]
__file__ = _os.path.join(_real, "__init__.py")
with open(__file__) as _f:
    exec(compile(_f.read(), __file__, "exec"))
del _os, _f, _real
