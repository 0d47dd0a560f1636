# mirrors models/ops/functions/__init__.py:9.  ms_deform_attn_core_pytorch keeps its name and
# signature but has no PyTorch/grid_sample body here: it runs the `border`-mode kernels (CUDA only);
# the grid_sample restatement lives in oracle/ as test infrastructure.
from .frontend import MSDeformFrontendFunction, msdeform_frontend
from .module_func import MSDeformModuleFunction, msdeform_module
from .ms_deform_attn_func import (MSDeformAttnBorderFunction, MSDeformAttnFunction, ms_deform_attn,
                                  ms_deform_attn_core_pytorch)

__all__ = ["MSDeformAttnFunction", "MSDeformAttnBorderFunction", "ms_deform_attn",
           "ms_deform_attn_core_pytorch", "MSDeformFrontendFunction", "msdeform_frontend",
           "MSDeformModuleFunction", "msdeform_module"]
