"""The module's front-end as one autograd Function (first step).

``msdeform_frontend(sampling_offsets, attention_logits, reference_points, level_sizes)``
returns ``(sampling_locations[N,Lq,M,L,P,2], attention_weights[N,Lq,M,L,P])`` -- the
softmax over L*P, ``ref + offsets / T_l`` (or the (centre, length) form) and the constant
``y = 0.5`` of models/ops/modules/ms_deform_attn.py:98-117 -- through
``msda_frontend_forward`` / ``msda_frontend_backward`` of libmsda_b200.so: one streaming
kernel each way instead of a softmax plus four or five elementwise kernels (and their
autograd twins).
"""
import torch
from torch.autograd import Function
from torch.autograd.function import once_differentiable

from .. import _lib
from ..MultiScaleDeformableAttention import _DTYPES, _stream


class MSDeformFrontendFunction(Function):
    @staticmethod
    def forward(ctx, sampling_offsets, attention_logits, reference_points, level_sizes):
        off, lg, ref = sampling_offsets, attention_logits, reference_points
        if not off.is_cuda:
            raise RuntimeError("Not implemented on the CPU")
        if off.dim() != 5 or ref.dim() != 4 or level_sizes.dim() != 1:
            raise RuntimeError("msda_b200 frontend: expected offsets[N,Lq,M,L,P], logits[N,Lq,M,L*P], "
                               "reference_points[N,Lq,L,1|2], level_sizes[L]")
        N, Lq, M, L, P = off.shape
        rd = ref.shape[-1]
        if rd not in (1, 2):
            raise ValueError("Last dim of reference_points must be 1 or 2, but get {} instead.".format(rd))
        if lg.numel() != off.numel() or tuple(ref.shape[:3]) != (N, Lq, L) or level_sizes.numel() != L:
            raise RuntimeError("msda_b200 frontend: inconsistent shapes offsets=%s logits=%s ref=%s sizes=%s" % (
                tuple(off.shape), tuple(lg.shape), tuple(ref.shape), tuple(level_sizes.shape)))
        if off.dtype not in _DTYPES or lg.dtype != off.dtype:
            raise RuntimeError("msda_b200 frontend: offsets and logits must share a dtype in "
                               "(float32, float64, bfloat16)")
        if level_sizes.dtype != torch.int64 or not level_sizes.is_cuda:
            raise RuntimeError("level_sizes must be a CUDA int64 tensor")
        off, lg = off.contiguous(), lg.contiguous()
        ref = ref.to(off.dtype).contiguous()
        level_sizes = level_sizes.contiguous()
        lib = _lib.load()
        with torch.cuda.device_of(off):
            loc = torch.empty((N, Lq, M, L, P, 2), dtype=off.dtype, device=off.device)
            aw = torch.empty((N, Lq, M, L, P), dtype=off.dtype, device=off.device)
            _lib.check(lib.msda_frontend_forward(
                off.data_ptr(), lg.data_ptr(), ref.data_ptr(), level_sizes.data_ptr(), loc.data_ptr(),
                aw.data_ptr(), N, Lq, M, L, P, rd, _DTYPES[off.dtype], _stream(off)))
        ctx.save_for_backward(off, aw, ref, level_sizes)
        ctx.logits_shape = attention_logits.shape
        ctx.ref_dtype = reference_points.dtype
        return loc, aw

    @staticmethod
    @once_differentiable
    def backward(ctx, grad_loc, grad_aw):
        off, aw, ref, level_sizes = ctx.saved_tensors
        N, Lq, M, L, P = off.shape
        rd = ref.shape[-1]
        lib = _lib.load()
        with torch.cuda.device_of(off):
            grad_loc = torch.zeros_like(off).new_zeros((N, Lq, M, L, P, 2)) if grad_loc is None \
                else grad_loc.to(off.dtype).contiguous()
            grad_aw = torch.zeros_like(aw) if grad_aw is None else grad_aw.to(off.dtype).contiguous()
            g_off = torch.empty_like(off)
            g_lg = torch.empty_like(aw)
            g_ref = torch.empty_like(ref) if ctx.needs_input_grad[2] else None
            _lib.check(lib.msda_frontend_backward(
                off.data_ptr(), aw.data_ptr(), ref.data_ptr(), level_sizes.data_ptr(), grad_loc.data_ptr(),
                grad_aw.data_ptr(), g_off.data_ptr(), g_lg.data_ptr(),
                g_ref.data_ptr() if g_ref is not None else None,
                N, Lq, M, L, P, rd, _DTYPES[off.dtype], _stream(off)))
        if g_ref is not None and g_ref.dtype != ctx.ref_dtype:
            g_ref = g_ref.to(ctx.ref_dtype)
        return g_off, g_lg.view(ctx.logits_shape), g_ref, None


def msdeform_frontend(sampling_offsets, attention_logits, reference_points, level_sizes):
    return MSDeformFrontendFunction.apply(sampling_offsets, attention_logits, reference_points, level_sizes)
