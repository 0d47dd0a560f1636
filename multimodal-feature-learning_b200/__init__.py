"""B200-native multi-scale deformable attention: the data-parallel hot path of
SAGA-DVC/multimodal-feature-learning (models/ops), rebuilt for sm_100a.

Layout mirrors the reference's models/ops package:
  functions.MSDeformAttnFunction   <- models/ops/functions/ms_deform_attn_func.py:23
  modules.MSDeformAttn             <- models/ops/modules/ms_deform_attn.py:30
  MultiScaleDeformableAttention    <- the pybind module (models/ops/src/vision.cpp)
"""
from . import _lib
from . import MultiScaleDeformableAttention
from .functions import MSDeformAttnFunction, ms_deform_attn
from .modules import MSDeformAttn

__all__ = ["MSDeformAttnFunction", "MSDeformAttn", "ms_deform_attn",
           "MultiScaleDeformableAttention", "_lib"]
