"""prepare_encoder_inputs (csrc/msda_prepare.cuh) against the reference method's own statements
(models/deformable/unimodal_deformable_transformer.py:104-134, restated below and run on the same GPU):
all six results bit-exact, gradients of the features, the positional embeddings and the level
embedding."""
import pytest
import torch

pytestmark = pytest.mark.gpu

import mfl_b200  # noqa: E402,F401
from mfl_b200.utils import prepare_encoder_inputs  # noqa: E402

DEV = "cuda"


def ref_prepare(srcs, masks, pos_embeds, level_embed):
    src_flatten, mask_flatten, lvl_pos_embed_flatten, temporal_shapes = [], [], [], []
    for lvl, (src, mask, pos_embed) in enumerate(zip(srcs, masks, pos_embeds)):
        temporal_shapes.append(src.shape[2])
        src = src.transpose(1, 2)
        pos_embed = pos_embed.transpose(1, 2)
        lvl_pos_embed_flatten.append(pos_embed + level_embed[lvl].view(1, 1, -1))
        src_flatten.append(src)
        mask_flatten.append(mask)
    src_flatten = torch.cat(src_flatten, 1)
    mask_flatten = torch.cat(mask_flatten, 1)
    lvl_pos_embed_flatten = torch.cat(lvl_pos_embed_flatten, 1)
    temporal_shapes = torch.as_tensor(temporal_shapes, dtype=torch.long, device=src_flatten.device)
    level_start_index = torch.cat((temporal_shapes.new_zeros((1,)), temporal_shapes.cumsum(0)[:-1]))
    valid_ratios = torch.stack([torch.sum(~m, 1).float() / m.shape[1] for m in masks], 1)      # get_valid_ratio, :86
    return src_flatten, temporal_shapes, level_start_index, valid_ratios, lvl_pos_embed_flatten, mask_flatten


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("sizes,C", [([300, 150, 75, 38], 512), ([33, 7], 48)])
def test_prepare_encoder_inputs(dtype, sizes, C):
    g = torch.Generator(device=DEV).manual_seed(9)
    B = 3
    srcs = [torch.randn(B, C, t, device=DEV, generator=g).to(dtype).requires_grad_(True) for t in sizes]
    poss = [torch.randn(B, C, t, device=DEV, generator=g).to(dtype).requires_grad_(True) for t in sizes]
    masks = [torch.rand(B, t, device=DEV, generator=g) > 0.8 for t in sizes]
    le = torch.randn(len(sizes), C, device=DEV, generator=g).to(dtype).requires_grad_(True)
    got = prepare_encoder_inputs(srcs, masks, poss, le)
    want = ref_prepare(srcs, masks, poss, le)
    for a, b in zip(got, want):
        assert a.shape == b.shape and a.dtype == b.dtype and torch.equal(a, b)
    w1, w2 = torch.randn_like(got[0]), torch.randn_like(got[4])
    ins = srcs + poss + [le]
    ga = torch.autograd.grad((got[0].float() * w1.float()).sum() + (got[4].float() * w2.float()).sum(), ins)
    gb = torch.autograd.grad((want[0].float() * w1.float()).sum() + (want[4].float() * w2.float()).sum(), ins)
    for a, b in zip(ga[:-1], gb[:-1]):
        assert torch.equal(a, b)
    tol = 1e-5 if dtype == torch.float32 else 2e-2
    assert float((ga[-1].float() - gb[-1].float()).abs().max()) <= tol * float(gb[-1].float().abs().max())
