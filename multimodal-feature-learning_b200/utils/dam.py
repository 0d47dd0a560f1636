"""Decoder attention map helpers -- drop-in for utils/dam.py of the reference
(consumer: models/criterion.py:246-303, the sparse models' mask-prediction loss).

``attn_map_to_flat_grid`` keeps the reference signature and result
(utils/dam.py:20-73) but is ONE kernel (``msda_dam_scatter`` in libmsda_b200.so)
instead of two permutes, two scatter_add_ passes and a dozen elementwise kernels.
No gradient flows through it in the reference either (its output only feeds a top-k).
"""
import torch

from .. import _lib
from ..MultiScaleDeformableAttention import _DTYPES, _stream


def idx_to_flat_grid(temporal_shapes, idx):
    """One-hot rows over the flat token grid (utils/dam.py:12-17)."""
    grid = torch.zeros((idx.shape[0], int(torch.sum(temporal_shapes))), device=idx.device, dtype=torch.float32)
    grid.scatter_(1, idx.to(torch.int64), 1)
    return grid


def attn_map_to_flat_grid(temporal_shapes, level_start_index, sampling_locations, attention_weights):
    """
    :param temporal_shapes       [L] int64, tokens per level
    :param level_start_index     [L] int64
    :param sampling_locations    [B, layers, Q, M, L, P, 1]  (or [..., 2] = the op's (x, y) pairs)
    :param attention_weights     [B, layers, Q, M, L, P]
    :return                      [B, layers, M, sum(temporal_shapes)] float32
    """
    loc, aw = sampling_locations, attention_weights
    if not loc.is_cuda:
        raise RuntimeError("Not implemented on the CPU")
    if loc.dim() != 7 or aw.dim() != 6 or tuple(loc.shape[:6]) != tuple(aw.shape) or loc.shape[6] not in (1, 2):
        raise RuntimeError("msda_b200 dam: expected sampling_locations[B,layers,Q,M,L,P,1|2] and "
                           "attention_weights[B,layers,Q,M,L,P], got %s and %s" % (tuple(loc.shape), tuple(aw.shape)))
    if loc.dtype not in _DTYPES or aw.dtype != loc.dtype:
        raise RuntimeError("msda_b200 dam: locations and weights must share a dtype in (float32, float64, bfloat16)")
    B, NL, Q, M, L, P, C = loc.shape
    sizes = temporal_shapes.to(device=loc.device, dtype=torch.int64).contiguous()
    start = level_start_index.to(device=loc.device, dtype=torch.int64).contiguous()
    if sizes.numel() != L or start.numel() != L:
        raise RuntimeError("msda_b200 dam: temporal_shapes / level_start_index must have one entry per level")
    S = int(sizes.sum())                       # the reference syncs here too (dam.py:53)
    loc, aw = loc.contiguous(), aw.contiguous()
    lib = _lib.load()
    with torch.cuda.device_of(loc):
        out = torch.empty((B, NL, M, S), dtype=torch.float32, device=loc.device)
        _lib.check(lib.msda_dam_scatter(loc.data_ptr(), aw.data_ptr(), sizes.data_ptr(), start.data_ptr(),
                                        out.data_ptr(), B * NL, Q, M, L, P, S, C, _DTYPES[loc.dtype], _stream(loc)))
    return out


def compute_corr(flat_grid_topk, flat_grid_attn_map, temporal_shapes):
    """Share of the attention mass that falls on the selected tokens, overall and per level
    (utils/dam.py:76-94)."""
    if flat_grid_topk.dim() == 1:
        flat_grid_topk, flat_grid_attn_map = flat_grid_topk[None], flat_grid_attn_map[None]
    hit = flat_grid_topk * flat_grid_attn_map
    corr = [hit.sum(-1) / flat_grid_attn_map.sum(-1)]
    lo = 0
    for t in temporal_shapes:
        hi = lo + int(t)
        corr.append(hit[:, lo:hi].sum(-1) / flat_grid_attn_map[:, lo:hi].sum(-1))
        lo = hi
    return corr
