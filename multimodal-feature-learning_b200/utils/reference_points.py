"""Encoder reference points -- drop-in for the ``get_reference_points`` static method of the
reference's encoders (models/deformable/unimodal_deformable_transformer.py:259-276; the same
code in multimodal_deformable_transformer.py:280-291 and both sparse encoders), as one kernel
(``msda_encoder_reference_points``) instead of a Python loop over a device tensor."""
import torch

from .. import _lib
from ..MultiScaleDeformableAttention import _stream


def get_reference_points(temporal_shapes, valid_ratios, device=None, num_tokens=None):
    """
    :param temporal_shapes  [L] int64, tokens per level
    :param valid_ratios     [N, L] float32
    :param device           kept for signature compatibility (the result lives with valid_ratios)
    :param num_tokens       sum(temporal_shapes) if the caller knows it (saves the host sync)
    :return                 [N, sum_l T_l, L, 1] float32
    """
    vr = valid_ratios
    if not vr.is_cuda:
        raise RuntimeError("Not implemented on the CPU")
    if vr.dim() != 2 or temporal_shapes.dim() != 1 or temporal_shapes.numel() != vr.shape[1]:
        raise RuntimeError("msda_b200: expected temporal_shapes[L] and valid_ratios[N, L]")
    N, L = vr.shape
    sizes = temporal_shapes.to(device=vr.device, dtype=torch.int64).contiguous()
    S = int(num_tokens) if num_tokens is not None else int(sizes.sum())
    vr32 = vr.to(torch.float32).contiguous()
    lib = _lib.load()
    with torch.cuda.device_of(vr):
        out = torch.empty((N, S, L, 1), dtype=torch.float32, device=vr.device)
        _lib.check(lib.msda_encoder_reference_points(sizes.data_ptr(), vr32.data_ptr(), out.data_ptr(), N, S, L,
                                                     _stream(vr)))
    return out
