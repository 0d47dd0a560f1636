"""MSDeformAttn nn.Module -- drop-in for models/ops/modules/ms_deform_attn.py:30-126
and for the live twin models/modules/attention.py:394-511 (extra ``is_sparse``
kwarg, xavier-initialised attention weights).

Same constructor, same parameter names/shapes (so reference checkpoints load:
sampling_offsets.*, attention_weights.*, value_proj.*, output_proj.*), same
forward signature.  The temporal (1-D) convention of this reference is kept:
``input_spatial_shapes`` is [L] (T_l) and sampling locations are scalar times;
the module stacks the constant y = 0.5 and shapes (1, T_l) before calling the
op (ms_deform_attn.py:114-117).  The core always runs on the sm_100a kernels;
a non-CUDA input raises (no PyTorch fallback).
"""
import math
import warnings
import weakref

import torch
import torch.nn.functional as F
from torch import nn
from torch.nn.init import constant_, xavier_uniform_

from ..functions import module_func
from ..functions.frontend import msdeform_frontend
from ..functions.ms_deform_attn_func import ms_deform_attn


def _is_power_of_2(n):
    if (not isinstance(n, int)) or (n < 0):
        raise ValueError("invalid input for _is_power_of_2: {} (type: {})".format(n, type(n)))
    return (n & (n - 1) == 0) and n != 0


class MSDeformAttn(nn.Module):
    def __init__(self, d_model=256, n_levels=4, n_heads=8, n_points=4, padding_mode="zeros",
                 attn_weight_init="zeros"):
        """
        :param d_model, n_levels, n_heads, n_points   as the reference (ms_deform_attn.py:31-38)
        :param padding_mode      'zeros'  = the reference CUDA op (what ms_deform_attn.py:119-122 runs on GPU)
                                 'border' = the reference PyTorch path (what models/modules/attention.py:505 runs)
        :param attn_weight_init  'zeros' (ms_deform_attn.py:72-73) or 'xavier' (attention.py:436-437)
        """
        super().__init__()
        if d_model % n_heads != 0:
            raise ValueError("d_model must be divisible by n_heads, but got {} and {}".format(d_model, n_heads))
        if padding_mode not in ("zeros", "border"):
            raise ValueError("padding_mode must be 'zeros' or 'border'")
        if not _is_power_of_2(d_model // n_heads):
            warnings.warn("You'd better set d_model in MSDeformAttn to make the dimension of each attention "
                          "head a power of 2 which is more efficient in our CUDA implementation.")
        self.im2col_step = 64  # ms_deform_attn.py:48
        self.d_model = d_model
        self.n_levels = n_levels
        self.n_heads = n_heads
        self.n_points = n_points
        self.padding_mode = padding_mode
        self.attn_weight_init = attn_weight_init
        self.check_input_lengths = True  # the reference's `assert shapes.sum() == Len_in` (a host sync)
        # softmax + sampling locations in one kernel (functions/frontend.py); False = the
        # reference's chain of torch kernels (ms_deform_attn.py:98-117), same results
        self.fused_frontend = True
        # ... and, where the call qualifies (functions/module_func.py: encoder-class shapes), that
        # arithmetic inside the op's tile kernels: no sampling_locations / attention_weights in memory
        self.fused_op = True
        self._len_cache = {}
        self._hw_cache = None   # (weakref to input_spatial_shapes, its version, the [L, 2] (1, T_l) tensor)

        self.sampling_offsets = nn.Linear(d_model, n_heads * n_levels * n_points)
        self.attention_weights = nn.Linear(d_model, n_heads * n_levels * n_points)
        self.value_proj = nn.Linear(d_model, d_model)
        self.output_proj = nn.Linear(d_model, d_model)
        self._reset_parameters()

    def __getstate__(self):
        # the caches hold a weakref / per-process verdicts: not part of a pickled module
        st = self.__dict__.copy()
        st["_hw_cache"] = None
        st["_len_cache"] = {}
        return st

    def _reset_parameters(self):
        # ms_deform_attn.py:62-77: offsets start on a ring of head directions,
        # point p at distance p+1; only the x component survives in 1-D use.
        constant_(self.sampling_offsets.weight.data, 0.)
        ang = torch.arange(self.n_heads, dtype=torch.float32) * (2.0 * math.pi / self.n_heads)
        ring = torch.stack([ang.cos(), ang.sin()], -1)
        ring_x = (ring / ring.abs().max(-1, keepdim=True)[0])[:, 0]
        scale = torch.arange(1, self.n_points + 1, dtype=torch.float32)
        bias = ring_x[:, None, None] * scale[None, None, :]
        bias = bias.expand(self.n_heads, self.n_levels, self.n_points)
        with torch.no_grad():
            self.sampling_offsets.bias = nn.Parameter(bias.reshape(-1).clone())
        if self.attn_weight_init == "xavier":
            xavier_uniform_(self.attention_weights.weight.data)
        else:
            constant_(self.attention_weights.weight.data, 0.)
        constant_(self.attention_weights.bias.data, 0.)
        xavier_uniform_(self.value_proj.weight.data)
        constant_(self.value_proj.bias.data, 0.)
        xavier_uniform_(self.output_proj.weight.data)
        constant_(self.output_proj.bias.data, 0.)

    def _shapes_hw(self, input_spatial_shapes):
        """[T_l] -> [(1, T_l)] (ms_deform_attn.py:116), built once per shapes tensor: the transformer
        passes the same tensor object to every layer and step."""
        c = self._hw_cache
        if c is not None and c[0]() is input_spatial_shapes and c[1] == input_spatial_shapes._version:
            return c[2]
        hw = torch.stack((torch.ones_like(input_spatial_shapes), input_spatial_shapes), -1)
        if not torch.cuda.is_current_stream_capturing():
            self._hw_cache = (weakref.ref(input_spatial_shapes), input_spatial_shapes._version, hw)
        return hw

    def _check_lengths(self, input_spatial_shapes, Len_in):
        # ms_deform_attn.py:93 syncs the device every call; cache the verdict per tensor OBJECT
        # (weak reference: a recycled address is a different tensor) and version, so steady-state
        # calls stay async.
        c = self._len_cache.get("last")
        if not (c is not None and c[0]() is input_spatial_shapes and c[1] == input_spatial_shapes._version
                and c[2] == Len_in):
            c = (weakref.ref(input_spatial_shapes), input_spatial_shapes._version, Len_in,
                 int(input_spatial_shapes.sum()) == Len_in)
            self._len_cache["last"] = c
        assert c[3], "sum of input_spatial_shapes must equal the flattened input length"

    def forward(self, query, reference_points, input_flatten, input_spatial_shapes,
                input_level_start_index, input_padding_mask=None, is_sparse=False):
        """
        :param query                    (N, Length_{query}, C)
        :param reference_points         (N, Length_{query}, n_levels, 1) in [0, 1], or (..., 2) = (centre, length)
        :param input_flatten            (N, sum_l T_l, C)
        :param input_spatial_shapes     (n_levels,) int64 [T_0 .. T_{L-1}]
        :param input_level_start_index  (n_levels,) int64 [0, T_0, T_0+T_1, ...]
        :param input_padding_mask       (N, sum_l T_l) bool, True = padding
        :param is_sparse                also return (sampling_locations[...,None], attention_weights)  (attention.py:508-511)
        :return output                  (N, Length_{query}, C)
        """
        N, Len_q, _ = query.shape
        N, Len_in, _ = input_flatten.shape
        if query.device.type != "cuda":
            raise RuntimeError("Not implemented on the CPU")  # models/ops/src/ms_deform_attn.h:38
        if self.check_input_lengths:
            self._check_lengths(input_spatial_shapes, Len_in)
        M, L, P = self.n_heads, self.n_levels, self.n_points

        value = self.value_proj(input_flatten).view(N, Len_in, M, self.d_model // M)
        sampling_offsets = self.sampling_offsets(query).view(N, Len_q, M, L, P)
        logits = self.attention_weights(query).view(N, Len_q, M, L * P)
        if reference_points.shape[-1] not in (1, 2):
            raise ValueError("Last dim of reference_points must be 1 or 2, but get {} instead.".format(
                reference_points.shape[-1]))
        if (self.fused_frontend and self.fused_op and not is_sparse
                and module_func.supported(value, sampling_offsets, reference_points)):
            # softmax + sampling locations inside the tile kernels: no loc / aw in memory
            # (the padding mask too: masked rows are zeroed in the kernels' value tile)
            output = module_func.msdeform_module(value, self._shapes_hw(input_spatial_shapes),
                                                 input_level_start_index, sampling_offsets, logits,
                                                 reference_points, self.padding_mode, input_padding_mask)
            return self.output_proj(output)
        if input_padding_mask is not None:
            value = value.masked_fill(input_padding_mask[..., None, None], float(0))
        if self.fused_frontend:
            loc_xy, attention_weights = msdeform_frontend(sampling_offsets, logits, reference_points,
                                                          input_spatial_shapes)
            sampling_locations = loc_xy[..., 0]
        else:
            attention_weights = F.softmax(logits, -1).view(N, Len_q, M, L, P)
            ref_t = reference_points[:, :, None, :, None, 0]
            if reference_points.shape[-1] == 1:
                sampling_locations = ref_t + sampling_offsets / input_spatial_shapes[None, None, None, :, None]
            else:
                sampling_locations = ref_t + sampling_offsets / P * reference_points[:, :, None, :, None, 1] * 0.5
            # temporal -> (x = t, y = 0.5) on maps of shape (1, T_l)   (ms_deform_attn.py:114-117)
            loc_xy = torch.stack((sampling_locations, torch.full_like(sampling_locations, 0.5)), -1)
        shapes_hw = self._shapes_hw(input_spatial_shapes)
        output = ms_deform_attn(value, shapes_hw, input_level_start_index, loc_xy.contiguous(),
                                attention_weights.contiguous(), self.im2col_step, self.padding_mode)
        output = self.output_proj(output)
        if is_sparse:
            return output, sampling_locations.unsqueeze(-1), attention_weights
        return output
