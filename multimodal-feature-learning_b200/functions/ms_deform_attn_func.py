"""Autograd surface of the op -- drop-in for
models/ops/functions/ms_deform_attn_func.py:23-41 (twin: models/modules/attention.py:311-328).

``MSDeformAttnFunction.apply(value, value_spatial_shapes, value_level_start_index,
sampling_locations, attention_weights, im2col_step)`` keeps the reference's
positional signature, output shape [N, Lq, M*D] and backward return pattern
``(grad_value, None, None, grad_sampling_loc, grad_attn_weight, None)``.
"""
import torch
from torch.autograd import Function
from torch.autograd.function import once_differentiable

from .. import MultiScaleDeformableAttention as MSDA
from .. import _lib


# Zero-fill of grad_value off the critical path: when `value` needs a gradient, its (f32/f64)
# grad_value buffer is allocated at FORWARD time and cleared on a side stream while the forward
# kernels run (they use a quarter of the DRAM bandwidth); the backward then accumulates into it
# without the 12 us memset in front (MSDA_FLAG_SKIP_ZERO_FILL).  Costs one value-sized buffer held
# between forward and backward; set to False to allocate and clear inside the backward instead.
PREZERO_GRAD_VALUE = True
_side_streams = {}


def _side_stream(device):
    key = (device.type, device.index if device.index is not None else torch.cuda.current_device())
    if key not in _side_streams:
        _side_streams[key] = torch.cuda.Stream(device=device)
    return _side_streams[key]


class _MSDeformAttnBase(Function):
    """Shared forward/backward; subclasses fix the padding mode so that
    ``apply`` keeps exactly six positional arguments."""

    padding_mode = "zeros"

    @classmethod
    def _fwd(cls, ctx, value, value_spatial_shapes, value_level_start_index,
             sampling_locations, attention_weights, im2col_step):
        ctx.im2col_step = im2col_step
        ctx.msda_gv = None
        if (PREZERO_GRAD_VALUE and value.requires_grad and value.is_cuda and value.numel() > 0
                and value.dtype in (torch.float32, torch.float64)):
            cur = torch.cuda.current_stream(value.device)
            side = _side_stream(value.device)
            gv = torch.empty_like(value)
            side.wait_stream(cur)            # the block may have been freed by work still queued on `cur`
            with torch.cuda.stream(side):
                gv.zero_()
            gv.record_stream(side)
            ev = torch.cuda.Event()
            ev.record(side)
            ctx.msda_gv = (gv, ev)
        output, plan = MSDA.ms_deform_attn_forward(
            value, value_spatial_shapes, value_level_start_index, sampling_locations,
            attention_weights, ctx.im2col_step, padding_mode=cls.padding_mode, return_plan=True,
            save_plan=MSDA.SAVE_PLAN and any(ctx.needs_input_grad))
        ctx.save_for_backward(value, value_spatial_shapes, value_level_start_index,
                              sampling_locations, attention_weights)
        ctx.msda_plan = plan     # time order of the queries (+ the tiles' saved plans when a backward follows)
        return output

    @classmethod
    def _bwd(cls, ctx, grad_output):
        value, shapes, lsi, loc, aw = ctx.saved_tensors
        # torch.use_deterministic_algorithms(True) selects the order-independent accumulation of
        # grad_value (the reference's atomicAdd scatter has no such mode)
        flags = _lib.MSDA_FLAG_DETERMINISTIC if torch.are_deterministic_algorithms_enabled() else 0
        gv = None
        if getattr(ctx, "msda_gv", None) is not None:
            gv, ev = ctx.msda_gv
            ctx.msda_gv = None          # single use: a second backward over the same graph (gradcheck,
                                        # retain_graph) allocates and clears its own buffer
            torch.cuda.current_stream(value.device).wait_event(ev)
            flags |= _lib.MSDA_FLAG_SKIP_ZERO_FILL      # (ignored by the deterministic variant)
        grad_value, grad_loc, grad_aw = MSDA.ms_deform_attn_backward(
            value, shapes, lsi, loc, aw, grad_output.contiguous(), ctx.im2col_step,
            padding_mode=cls.padding_mode, flags=flags, plan=getattr(ctx, "msda_plan", None),
            grad_value=gv)
        return grad_value, None, None, grad_loc, grad_aw, None


class MSDeformAttnFunction(_MSDeformAttnBase):
    """Zero padding outside the map == the reference CUDA op
    (ms_deform_im2col_cuda.cuh:56-79, :289)."""

    padding_mode = "zeros"

    @staticmethod
    def forward(ctx, value, value_spatial_shapes, value_level_start_index, sampling_locations,
                attention_weights, im2col_step):
        return MSDeformAttnFunction._fwd(ctx, value, value_spatial_shapes, value_level_start_index,
                                         sampling_locations, attention_weights, im2col_step)

    @staticmethod
    @once_differentiable
    def backward(ctx, grad_output):
        return MSDeformAttnFunction._bwd(ctx, grad_output)


class MSDeformAttnBorderFunction(_MSDeformAttnBase):
    """Border padding == the reference's PyTorch path, which every transformer
    in the reference actually runs (grid_sample padding_mode='border',
    ms_deform_attn_func.py:61-62 / models/modules/attention.py:367-368)."""

    padding_mode = "border"

    @staticmethod
    def forward(ctx, value, value_spatial_shapes, value_level_start_index, sampling_locations,
                attention_weights, im2col_step):
        return MSDeformAttnBorderFunction._fwd(ctx, value, value_spatial_shapes,
                                               value_level_start_index, sampling_locations,
                                               attention_weights, im2col_step)

    @staticmethod
    @once_differentiable
    def backward(ctx, grad_output):
        return MSDeformAttnBorderFunction._bwd(ctx, grad_output)


def ms_deform_attn(value, value_spatial_shapes, value_level_start_index, sampling_locations,
                   attention_weights, im2col_step=64, padding_mode="zeros"):
    """Functional form with a selectable padding mode."""
    fn = {"zeros": MSDeformAttnFunction, "border": MSDeformAttnBorderFunction}[padding_mode]
    return fn.apply(value, value_spatial_shapes, value_level_start_index, sampling_locations,
                    attention_weights, im2col_step)


def ms_deform_attn_core_pytorch(value, value_spatial_shapes, sampling_locations, attention_weights,
                                return_value=False):
    """The name the reference exports next to the autograd Function
    (models/ops/functions/__init__.py:9; imported by models/ops/test.py:18 and
    models/ops/modules/ms_deform_attn.py:21).  There it is the grid_sample (`border` padding)
    restatement of the op; here the same function -- same arguments, same [N, Lq, M*D] result,
    differentiable -- runs the `border`-mode kernels (MSDeformAttnBorderFunction).  CUDA tensors
    only: CPU tensors raise "Not implemented on the CPU" like everything else in this package.
    `return_value=True` (the per-point sampled values, MSDeformAttnCap) is out of scope."""
    if return_value:
        raise NotImplementedError("return_value=True (MSDeformAttnCap) is outside this package's hot path")
    shapes = torch.as_tensor(value_spatial_shapes, dtype=torch.long, device=value.device).reshape(-1, 2)
    sizes = shapes[:, 0] * shapes[:, 1]
    lsi = torch.cat((sizes.new_zeros((1,)), sizes.cumsum(0)[:-1]))
    return MSDeformAttnBorderFunction.apply(value.contiguous(), shapes.contiguous(), lsi,
                                            sampling_locations.contiguous(),
                                            attention_weights.contiguous(), 64)
