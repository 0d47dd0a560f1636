"""The module's front-end inside the op (second step).

``msdeform_module(value, shapes_hw, level_start_index, sampling_offsets, attention_logits,
reference_points, padding_mode)`` is ``ms_deform_attn(value, ..., *msdeform_frontend(offsets, logits,
ref, T), ...)`` -- models/ops/modules/ms_deform_attn.py:98-122 from the two Linear outputs to the
op's output -- as ONE call each way: the CTA-tile kernels do the softmax and the location
arithmetic in their prologue (``msda_module_forward`` / ``msda_module_backward``), so
``sampling_locations`` and ``attention_weights`` are never materialised.  ``supported(...)`` tells
whether a call qualifies (temporal maps, P = 4, L <= 4, Lq >= 256, f32 / bf16, D = 64 or 32);
MSDeformAttn falls back to the two-kernel path otherwise (and for ``is_sparse=True``, which must
return the locations and weights)."""
import torch
from torch.autograd import Function
from torch.autograd.function import once_differentiable

from .. import _lib
from .. import MultiScaleDeformableAttention as MSDA
from ..MultiScaleDeformableAttention import _DTYPES, _stream
from . import ms_deform_attn_func as _f


def supported(value, offsets, reference_points, flags=0):
    if not value.is_cuda or value.dtype not in (torch.float32, torch.bfloat16):
        return False
    if torch.are_deterministic_algorithms_enabled():
        return False
    N, S, M, D = value.shape
    _, Lq, _, L, P = offsets.shape
    return bool(_lib.load().msda_module_supported(N, S, M, D, L, Lq, P, reference_points.shape[-1],
                                                  _DTYPES[value.dtype], flags))


class MSDeformModuleFunction(Function):
    @staticmethod
    def forward(ctx, value, shapes_hw, lsi, offsets, logits, ref, padding_mode, mask=None):
        N, S, M, D = value.shape
        _, Lq, _, L, P = offsets.shape
        rd = ref.shape[-1]
        dt = _DTYPES[value.dtype]
        value, offsets, logits = value.contiguous(), offsets.contiguous(), logits.contiguous()
        ref_c = ref.to(value.dtype).contiguous()
        if mask is not None:
            if mask.dtype != torch.bool or tuple(mask.shape) != (N, S):
                raise RuntimeError("padding mask must be a bool [N, S] tensor")
            mask = mask.contiguous()
        lib = _lib.load()
        ctx.msda_gv = None
        with torch.cuda.device_of(value):
            if _f.PREZERO_GRAD_VALUE and value.requires_grad and value.dtype == torch.float32:
                cur = torch.cuda.current_stream(value.device)
                side = _f._side_stream(value.device)
                gv = torch.empty_like(value)
                side.wait_stream(cur)
                with torch.cuda.stream(side):
                    gv.zero_()
                gv.record_stream(side)
                ev = torch.cuda.Event()
                ev.record(side)
                ctx.msda_gv = (gv, ev)
            out = torch.empty((N, Lq, M * D), dtype=value.dtype, device=value.device)
            # the tiles' plans travel to the backward with the query order when there will be one
            fflags = _lib.MSDA_FLAG_SAVE_PLAN if (MSDA.SAVE_PLAN and any(ctx.needs_input_grad)
                                                  and MSDA.saved_plans_supported()) else 0
            ctx.msda_fflags = fflags
            ws_bytes = lib.msda_forward_workspace_bytes(N, S, M, D, L, Lq, P, dt, fflags)
            plan = torch.empty(max(ws_bytes // 4, 1), dtype=torch.int32, device=value.device)
            _lib.check(lib.msda_module_forward(
                value.data_ptr(), shapes_hw.data_ptr(), lsi.data_ptr(), offsets.data_ptr(), logits.data_ptr(),
                ref_c.data_ptr(), mask.data_ptr() if mask is not None else None, out.data_ptr(),
                N, S, M, D, L, Lq, P, rd, dt,
                _lib.PADDING_MODES[padding_mode], fflags, plan.data_ptr(), ws_bytes, _stream(value)))
        ctx.save_for_backward(value, shapes_hw, lsi, offsets, logits, ref_c, plan,
                              mask if mask is not None else plan.new_empty(0))
        ctx.has_mask = mask is not None
        ctx.padding_mode, ctx.logits_shape, ctx.ref_dtype = padding_mode, logits.shape, ref.dtype
        return out

    @staticmethod
    @once_differentiable
    def backward(ctx, grad_out):
        value, shapes_hw, lsi, offsets, logits, ref, plan, mask = ctx.saved_tensors
        N, S, M, D = value.shape
        _, Lq, _, L, P = offsets.shape
        rd = ref.shape[-1]
        dt = _DTYPES[value.dtype]
        lib = _lib.load()
        flags = ctx.msda_fflags
        with torch.cuda.device_of(value):
            gv = None
            if ctx.msda_gv is not None:
                gv, ev = ctx.msda_gv
                ctx.msda_gv = None
                torch.cuda.current_stream(value.device).wait_event(ev)
                flags |= _lib.MSDA_FLAG_SKIP_ZERO_FILL
            if gv is None:
                gv = torch.empty_like(value)
            g_off = torch.empty_like(offsets)
            g_lg = torch.empty_like(logits)
            g_ref = torch.empty(ref.shape, dtype=torch.float32, device=value.device) if ctx.needs_input_grad[5] else None
            ws = None
            ws_bytes = N * S * M * D * 4 if value.dtype == torch.bfloat16 else 0
            if ws_bytes:
                ws = torch.empty(ws_bytes // 4, dtype=torch.float32, device=value.device)
            _lib.check(lib.msda_module_backward(
                value.data_ptr(), shapes_hw.data_ptr(), lsi.data_ptr(), offsets.data_ptr(), logits.data_ptr(),
                ref.data_ptr(), mask.data_ptr() if ctx.has_mask else None, grad_out.contiguous().data_ptr(),
                gv.data_ptr(), g_off.data_ptr(), g_lg.data_ptr(),
                g_ref.data_ptr() if g_ref is not None else None, N, S, M, D, L, Lq, P, rd, dt,
                _lib.PADDING_MODES[ctx.padding_mode], flags, ws.data_ptr() if ws is not None else None, ws_bytes,
                plan.data_ptr(), _stream(value)))
            if ws is not None:
                ws.record_stream(torch.cuda.current_stream(value.device))
        if g_ref is not None:
            g_ref = g_ref.to(ctx.ref_dtype)
        return gv, None, None, g_off, g_lg.view(ctx.logits_shape), g_ref, None, None


def msdeform_module(value, shapes_hw, level_start_index, sampling_offsets, attention_logits, reference_points,
                    padding_mode="zeros", padding_mask=None):
    """padding_mask [N, S] bool (True = padding): value.masked_fill(mask[..., None], 0) of
    ms_deform_attn.py:96-97, applied inside the kernels (pass the UNMASKED value)."""
    return MSDeformModuleFunction.apply(value, shapes_hw, level_start_index, sampling_offsets, attention_logits,
                                        reference_points, padding_mode, padding_mask)
