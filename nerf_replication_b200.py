"""Import alias: the package directory is named ``nerf-replication_b200`` (not a valid Python identifier),
so this module makes it importable as ``nerf_replication_b200`` by pointing ``__path__`` at it."""
import os as _os

__path__ = [_os.path.join(_os.path.dirname(_os.path.abspath(__file__)), "nerf-replication_b200")]
with open(_os.path.join(__path__[0], "__init__.py")) as _f:
    exec(compile(_f.read(), _os.path.join(__path__[0], "__init__.py"), "exec"))
del _os, _f
