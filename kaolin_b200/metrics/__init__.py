# mirrors kaolin/metrics/__init__.py for the module on the path's consumer side
from . import render  # noqa: F401
