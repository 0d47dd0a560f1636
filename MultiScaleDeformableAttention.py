"""Top-level stand-in for the reference's compiled extension: `import MultiScaleDeformableAttention
as MSDA` (models/ops/functions/ms_deform_attn_func.py:18-21, models/modules/attention.py:15-18)
resolves to this repo's C-ABI-backed module when the repo root is on sys.path, so the reference's
own Python files run unedited on top of libmsda_b200.so."""
import os
import sys

_root = os.path.dirname(os.path.abspath(__file__))
if _root not in sys.path:
    sys.path.insert(0, _root)
import mfl_b200  # noqa: E402

_impl = mfl_b200.MultiScaleDeformableAttention
ms_deform_attn_forward = _impl.ms_deform_attn_forward
ms_deform_attn_backward = _impl.ms_deform_attn_backward
