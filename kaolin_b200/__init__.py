"""kaolin_b200 — B200-native (sm_100a) implementation of Kaolin's DIB-R differentiable rasterization path.

Drop-in for ``kaolin.render.mesh.{rasterize, dibr_soft_mask, dibr_rasterization}`` and for the four native ops
``kaolin._C.render.mesh.*`` they sit on, plus the callers either side of the path:
``kaolin.render.mesh.{prepare_vertices, texture_mapping}`` and ``kaolin.metrics.render.mask_iou``.
Nothing else of Kaolin is provided here.
"""
from . import _C  # noqa: F401
from . import render  # noqa: F401
from . import metrics  # noqa: F401

__version__ = "0.1.0"
