"""TEST INFRASTRUCTURE -- port of the reference's pure-PyTorch core.

Restates ms_deform_attn_core_pytorch
(models/ops/functions/ms_deform_attn_func.py:44-71; 1-D twin at
models/modules/attention.py:331-383): per level, view the value slab as an
image batch (N*M, D, H, W), bilinearly sample it with F.grid_sample at
2*loc-1 (align_corners=False), weight by the attention weights and sum over
levels*points.  The reference hard-codes padding_mode='border'; here it is a
parameter so the same code also yields the zero-padded variant (= what the
reference CUDA kernel computes).  This is the CPU baseline bench.py times
(cpu_baseline.kind = "port") and one of the oracles the tests use.
"""
import torch
import torch.nn.functional as F


def ms_deform_attn_core_pytorch(value, value_spatial_shapes, sampling_locations, attention_weights,
                                padding_mode="border"):
    n, _, m, d = value.shape
    _, lq, _, n_lvl, n_pt, _ = sampling_locations.shape
    sizes = [(int(h), int(w)) for h, w in value_spatial_shapes.tolist()]
    slabs = value.split([h * w for h, w in sizes], dim=1)
    grid = sampling_locations * 2 - 1
    sampled = []
    for lvl, (h, w) in enumerate(sizes):
        img = slabs[lvl].permute(0, 2, 3, 1).reshape(n * m, d, h, w)
        g = grid[:, :, :, lvl].permute(0, 2, 1, 3, 4).reshape(n * m, lq, n_pt, 2)
        sampled.append(F.grid_sample(img, g, mode="bilinear", padding_mode=padding_mode,
                                     align_corners=False))          # (N*M, D, Lq, P)
    stacked = torch.cat(sampled, dim=-1)                             # (N*M, D, Lq, L*P)
    w = attention_weights.permute(0, 2, 1, 3, 4).reshape(n * m, 1, lq, n_lvl * n_pt)
    out = (stacked * w).sum(-1)                                      # (N*M, D, Lq)
    return out.view(n, m * d, lq).permute(0, 2, 1).contiguous()


def frontend_pytorch(sampling_offsets, attention_logits, reference_points, level_sizes):
    """Restatement of the module's front-end (models/ops/modules/ms_deform_attn.py:98-117;
    live twin models/modules/attention.py:462-496): softmax over L*P, sampling locations
    from reference points and offsets, the constant y = 0.5.  Oracle for
    msda_frontend_forward / _backward (gradients through autograd)."""
    n, lq, m, n_lvl, n_pt = sampling_offsets.shape
    aw = F.softmax(attention_logits.reshape(n, lq, m, n_lvl * n_pt), -1).view(n, lq, m, n_lvl, n_pt)
    ref0 = reference_points[:, :, None, :, None, 0]
    if reference_points.shape[-1] == 1:
        x = ref0 + sampling_offsets / level_sizes[None, None, None, :, None]
    elif reference_points.shape[-1] == 2:
        x = ref0 + sampling_offsets / n_pt * reference_points[:, :, None, :, None, 1] * 0.5
    else:
        raise ValueError("Last dim of reference_points must be 1 or 2")
    loc = torch.stack((x, 0.5 * x.new_ones(x.shape)), -1)
    return loc, aw


def reference_points_pytorch(temporal_shapes, valid_ratios):
    """Restatement of DeformableTransformerEncoder.get_reference_points
    (models/deformable/unimodal_deformable_transformer.py:259-276): token t of level k gets, for
    every level l, (t + 0.5) / (valid_ratio_k * T_k) * valid_ratio_l.  Oracle for
    msda_encoder_reference_points."""
    cols = []
    for k, t_k in enumerate(int(t) for t in temporal_shapes.tolist()):
        centre = torch.arange(t_k, dtype=torch.float32, device=valid_ratios.device) + 0.5
        cols.append(centre[None] / (valid_ratios[:, None, k] * t_k))
    ref = torch.cat(cols, 1)
    return (ref[:, :, None] * valid_ratios[:, None])[..., None]
