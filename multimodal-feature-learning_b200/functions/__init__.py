# mirrors models/ops/functions/__init__.py:9 (the PyTorch fallback
# ms_deform_attn_core_pytorch is NOT re-exported: this package has no CPU path;
# its restatement lives in oracle/ as test infrastructure only)
from .frontend import MSDeformFrontendFunction, msdeform_frontend
from .ms_deform_attn_func import MSDeformAttnFunction, ms_deform_attn

__all__ = ["MSDeformAttnFunction", "ms_deform_attn", "MSDeformFrontendFunction", "msdeform_frontend"]
