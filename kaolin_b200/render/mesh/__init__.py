# mirrors kaolin/render/mesh/__init__.py:1-4 for the hot-path modules
from .rasterization import *  # noqa: F401,F403
from .dibr import *  # noqa: F401,F403
from .dibr import dibr_rasterization_from_mesh  # noqa: F401  (not in the reference: prepare_vertices folded into K0)
from .utils import *  # noqa: F401,F403
from .deftet import *  # noqa: F401,F403
from . import rasterization, dibr, utils, deftet  # noqa: F401
