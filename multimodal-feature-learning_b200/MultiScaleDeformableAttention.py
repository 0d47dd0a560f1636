"""Drop-in for the reference's pybind module ``MultiScaleDeformableAttention``
(models/ops/src/vision.cpp:13-16, built by models/ops/setup.py:53).

Exports the same two functions with the same positional arguments and the same
error behaviour (asserts -> RuntimeError), but the work is done by the sm_100a
kernels in libmsda_b200.so through its C ABI (include/msda_b200.h).  The extra
keyword arguments (padding_mode, flags) are optional extensions.
"""
import torch

from . import _lib

_DTYPES = {torch.float32: _lib.MSDA_F32, torch.float64: _lib.MSDA_F64,
           torch.bfloat16: _lib.MSDA_BF16}


def _check_inputs(named, im2col_step):
    # models/ops/src/cuda/ms_deform_attn_cuda.cu:28-38 / :93-105
    value = named[0][1]
    if not value.is_cuda:
        # models/ops/src/ms_deform_attn.h:38,60
        raise RuntimeError("Not implemented on the CPU")
    for name, t in named:
        if not t.is_contiguous():
            raise RuntimeError("%s tensor has to be contiguous" % name)
        if not t.is_cuda:
            raise RuntimeError("%s must be a CUDA tensor" % name)
        if t.device != value.device:
            raise RuntimeError("%s must be on the same device as value" % name)
    value, shapes, lsi, loc, aw = (t for _, t in named[:5])
    if value.dtype not in _DTYPES:
        raise RuntimeError("msda_b200: unsupported dtype %s (float32, float64, bfloat16)" % value.dtype)
    for name, t in named[3:]:
        if t.dtype != value.dtype:
            raise RuntimeError("%s must have the dtype of value (%s), got %s" % (name, value.dtype, t.dtype))
    if shapes.dtype != torch.int64 or lsi.dtype != torch.int64:
        raise RuntimeError("spatial_shapes and level_start_index must be int64 tensors")
    if value.dim() != 4 or loc.dim() != 6 or aw.dim() != 5 or shapes.dim() != 2:
        raise RuntimeError("msda_b200: expected value[N,S,M,D], spatial_shapes[L,2], "
                           "sampling_loc[N,Lq,M,L,P,2], attn_weight[N,Lq,M,L,P]")
    N, S, M, D = value.shape
    L = shapes.shape[0]
    Lq, P = loc.shape[1], loc.shape[4]
    if (shapes.shape[1] != 2 or lsi.numel() != L or tuple(loc.shape) != (N, Lq, M, L, P, 2)
            or tuple(aw.shape) != (N, Lq, M, L, P)):
        raise RuntimeError("msda_b200: inconsistent shapes value=%s shapes=%s lsi=%s loc=%s aw=%s" % (
            tuple(value.shape), tuple(shapes.shape), tuple(lsi.shape), tuple(loc.shape), tuple(aw.shape)))
    # cuda.cu:50-52: im2col_step_ = min(batch, im2col_step); batch % im2col_step_ == 0
    if N > 0:
        step = min(N, int(im2col_step))
        if step <= 0 or N % step != 0:
            raise RuntimeError("batch(%d) must divide im2col_step(%d)" % (N, step))
    return N, S, M, D, L, Lq, P


def _stream(t):
    return torch.cuda.current_stream(t.device).cuda_stream


# True: the plan a forward returns for its backward also holds every tile's windows and merged
# weights (MSDA_FLAG_SAVE_PLAN, ~1.8 KB per query at M = 8) and the backward starts from them instead
# of the sampling points.  Same results; measured on B200 it does not pay (forward + 2 us for the
# stores, backward unchanged: DESIGN.md section 6), so only the query order (4 B per query) is kept;
# the default build of the library ignores the flag altogether (-DMSDA_TILE_SAVED_PLANS=1 enables it).
SAVE_PLAN = False


def saved_plans_supported():
    """Whether libmsda_b200.so was built with -DMSDA_TILE_SAVED_PLANS=1."""
    lib = _lib.load()
    a = lib.msda_forward_workspace_bytes(2, 1500, 8, 64, 4, 1500, 4, 0, _lib.MSDA_FLAG_SAVE_PLAN)
    return a != lib.msda_forward_workspace_bytes(2, 1500, 8, 64, 4, 1500, 4, 0, 0)


def ms_deform_attn_forward(value, spatial_shapes, level_start_index, sampling_loc, attn_weight,
                           im2col_step, padding_mode="zeros", flags=0, return_plan=False, save_plan=None):
    """value[N,S,M,D] ... -> output[N,Lq,M*D]  (cuda.cu:20-80).
    return_plan=True also returns the plan workspace of msda_forward_ex (the time order of the
    queries; None when the shape does not use the CTA-tile kernels), which
    ms_deform_attn_backward(plan=...) reuses instead of sorting again.  save_plan (default: the
    module switch SAVE_PLAN, when a plan is returned): the plan also carries the tiles' saved plans
    (plan.msda_saved is True then)"""
    N, S, M, D, L, Lq, P = _check_inputs(
        [("value", value), ("spatial_shapes", spatial_shapes),
         ("level_start_index", level_start_index), ("sampling_loc", sampling_loc),
         ("attn_weight", attn_weight)], im2col_step)
    lib = _lib.load()
    dt = _DTYPES[value.dtype]
    if save_plan is None:
        save_plan = SAVE_PLAN and return_plan
    if save_plan and not (flags & _lib.MSDA_FLAG_PLAN):
        flags |= _lib.MSDA_FLAG_SAVE_PLAN
    with torch.cuda.device_of(value):
        out = torch.empty((N, Lq, M * D), dtype=value.dtype, device=value.device)
        ws_bytes = lib.msda_forward_workspace_bytes(N, S, M, D, L, Lq, P, dt, flags)
        if (flags & _lib.MSDA_FLAG_SAVE_PLAN) and ws_bytes == lib.msda_forward_workspace_bytes(
                N, S, M, D, L, Lq, P, dt, flags & ~_lib.MSDA_FLAG_SAVE_PLAN):
            flags &= ~_lib.MSDA_FLAG_SAVE_PLAN     # this build of the library does not save plans
        ws = torch.empty(ws_bytes // 4, dtype=torch.int32, device=value.device) if ws_bytes else None
        _lib.check(lib.msda_forward_ex(
            value.data_ptr(), spatial_shapes.data_ptr(), level_start_index.data_ptr(),
            sampling_loc.data_ptr(), attn_weight.data_ptr(), out.data_ptr(),
            N, S, M, D, L, Lq, P, dt, _lib.PADDING_MODES[padding_mode],
            flags, ws.data_ptr() if ws is not None else None, ws_bytes, _stream(value)))
        if ws is not None:
            ws.record_stream(torch.cuda.current_stream(value.device))
            ws.msda_saved = bool(flags & _lib.MSDA_FLAG_SAVE_PLAN)
    return (out, ws) if return_plan else out


def ms_deform_attn_backward(value, spatial_shapes, level_start_index, sampling_loc, attn_weight,
                            grad_output, im2col_step, padding_mode="zeros", flags=0, plan=None,
                            grad_value=None):
    """-> [grad_value, grad_sampling_loc, grad_attn_weight]  (cuda.cu:83-153).
    flags |= MSDA_FLAG_DETERMINISTIC (16): bitwise reproducible grad_value (fixed-point
    accumulation, csrc/msda_det.cuh); the autograd Functions set it when
    torch.are_deterministic_algorithms_enabled().
    plan: the forward's plan workspace (same sampling_loc).  grad_value: an already ZEROED
    buffer to accumulate into (pass flags |= MSDA_FLAG_SKIP_ZERO_FILL with it; f32/f64 only)."""
    N, S, M, D, L, Lq, P = _check_inputs(
        [("value", value), ("spatial_shapes", spatial_shapes),
         ("level_start_index", level_start_index), ("sampling_loc", sampling_loc),
         ("attn_weight", attn_weight), ("grad_output", grad_output)], im2col_step)
    if grad_output.numel() != N * Lq * M * D:
        raise RuntimeError("grad_output must have N*Lq*M*D elements")
    lib = _lib.load()
    dt = _DTYPES[value.dtype]
    flags &= ~_lib.MSDA_FLAG_SAVE_PLAN
    if plan is not None and getattr(plan, "msda_saved", False):
        flags |= _lib.MSDA_FLAG_SAVE_PLAN      # the plan carries the forward's tile plans
    with torch.cuda.device_of(value):
        if grad_value is None:
            grad_value = torch.empty_like(value)
        grad_loc = torch.empty_like(sampling_loc)
        grad_aw = torch.empty_like(attn_weight)
        ws_bytes = lib.msda_backward_workspace_bytes_flags(N, S, M, D, L, Lq, P, dt, flags)
        ws = None
        if ws_bytes:
            ws = torch.empty((ws_bytes + 3) // 4, dtype=torch.float32, device=value.device)
        _lib.check(lib.msda_backward_ex(
            value.data_ptr(), spatial_shapes.data_ptr(), level_start_index.data_ptr(),
            sampling_loc.data_ptr(), attn_weight.data_ptr(), grad_output.data_ptr(),
            grad_value.data_ptr(), grad_loc.data_ptr(), grad_aw.data_ptr(),
            N, S, M, D, L, Lq, P, dt, _lib.PADDING_MODES[padding_mode], flags,
            ws.data_ptr() if ws is not None else None, ws_bytes,
            plan.data_ptr() if plan is not None else None, _stream(value)))
        if ws is not None:
            ws.record_stream(torch.cuda.current_stream(value.device))
    return [grad_value, grad_loc, grad_aw]
