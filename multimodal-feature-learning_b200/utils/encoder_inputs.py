"""``prepare_encoder_inputs`` of the deformable transformers
(models/deformable/unimodal_deformable_transformer.py:90-134 and its copies) as two launches:
``msda_prepare_encoder_inputs`` (second half).  Same arguments (plus the
level embedding, a module parameter there) and the same six results; differentiable w.r.t. the
features, the positional embeddings and the level embedding (the backward is views and one small
reduction per level).  CUDA tensors only."""
import ctypes

import torch
from torch.autograd import Function
from torch.autograd.function import once_differentiable

from .. import _lib
from ..MultiScaleDeformableAttention import _DTYPES, _stream


class _PrepareEncoderInputs(Function):
    @staticmethod
    def forward(ctx, level_embed, n_levels, *tensors):
        srcs, masks, poss = tensors[:n_levels], tensors[n_levels:2 * n_levels], tensors[2 * n_levels:]
        s0 = srcs[0]
        if not s0.is_cuda:
            raise RuntimeError("Not implemented on the CPU")
        B, C = s0.shape[0], s0.shape[1]
        sizes = [int(s.shape[2]) for s in srcs]
        S = sum(sizes)
        srcs = [s.contiguous() for s in srcs]
        poss = [p.to(s0.dtype).contiguous() for p in poss]
        masks = [m.contiguous() for m in masks]
        for s, p, m, t in zip(srcs, poss, masks, sizes):
            if tuple(s.shape) != (B, C, t) or tuple(p.shape) != (B, C, t) or tuple(m.shape) != (B, t) \
                    or m.dtype != torch.bool or s.dtype != s0.dtype:
                raise RuntimeError("prepare_encoder_inputs: srcs / pos_embeds [B, C, T_l] and masks [B, T_l] (bool) expected")
        L = n_levels
        arr = ctypes.c_void_p * L
        out_src = torch.empty(B, S, C, dtype=s0.dtype, device=s0.device)
        out_pos = torch.empty_like(out_src)
        out_mask = torch.empty(B, S, dtype=torch.bool, device=s0.device)
        vr = torch.empty(B, L, dtype=torch.float32, device=s0.device)
        host_sizes = (ctypes.c_int64 * L)(*sizes)
        with torch.cuda.device_of(s0):
            _lib.check(_lib.load().msda_prepare_encoder_inputs(
                arr(*[t.data_ptr() for t in srcs]), arr(*[t.data_ptr() for t in poss]),
                arr(*[t.data_ptr() for t in masks]), host_sizes, level_embed.to(s0.dtype).contiguous().data_ptr(),
                out_src.data_ptr(), out_pos.data_ptr(), out_mask.data_ptr(), vr.data_ptr(), B, C, L,
                _DTYPES[s0.dtype], _stream(s0)))
        ctx.sizes, ctx.n_levels, ctx.le_dtype = sizes, L, level_embed.dtype
        ctx.mark_non_differentiable(out_mask, vr)
        return out_src, out_pos, out_mask, vr

    @staticmethod
    @once_differentiable
    def backward(ctx, g_src, g_pos, _gm, _gv):
        L, sizes = ctx.n_levels, ctx.sizes
        gs = g_src.split(sizes, 1) if g_src is not None else [None] * L
        gp = g_pos.split(sizes, 1) if g_pos is not None else [None] * L
        g_le = None
        if g_pos is not None:
            g_le = torch.stack([x.sum((0, 1)) for x in gp]).to(ctx.le_dtype)
        tr = lambda x: None if x is None else x.transpose(1, 2)
        return (g_le, None) + tuple(tr(x) for x in gs) + (None,) * L + tuple(tr(x) for x in gp)


def prepare_encoder_inputs(srcs, masks, pos_embeds, level_embed):
    """-> (src_flatten [B,S,C], temporal_shapes [L] int64, level_start_index [L] int64,
    valid_ratios [B,L], lvl_pos_embed_flatten [B,S,C], mask_flatten [B,S]) -- the reference's order."""
    L = len(srcs)
    out_src, out_pos, out_mask, vr = _PrepareEncoderInputs.apply(level_embed, L, *srcs, *masks, *pos_embeds)
    sizes = [int(s.shape[2]) for s in srcs]
    temporal_shapes = torch.as_tensor(sizes, dtype=torch.long, device=out_src.device)
    starts = [0]
    for t in sizes[:-1]:
        starts.append(starts[-1] + t)
    level_start_index = torch.as_tensor(starts, dtype=torch.long, device=out_src.device)
    return out_src, temporal_shapes, level_start_index, vr, out_pos, out_mask
