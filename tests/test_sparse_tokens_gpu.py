"""Sparse-encoder token gather / scatter (csrc/msda_sparse.cuh) against the reference's own torch
expressions (models/sparse/unimodal_sparse_deformable_transformer.py:425-431 and :443-448, restated
below line by line and run on the same GPU), forward and gradients, with and without per-item
token counts, fp32 / bf16, a row size that is not a multiple of 16 bytes."""
import pytest
import torch

pytestmark = pytest.mark.gpu

import mfl_b200  # noqa: E402,F401
from mfl_b200.utils import gather_tokens, scatter_tokens  # noqa: E402

DEV = "cuda"


def ref_gather(x, idx):                      # :428-430
    return torch.gather(x, 1, idx.unsqueeze(-1).repeat(1, 1, x.size(-1)))


def ref_scatter(output, idx, tgt, nums):     # :441-448
    if nums is None:
        return output.scatter(1, idx.unsqueeze(-1).repeat(1, 1, tgt.size(-1)), tgt)
    outs = []
    for i in range(idx.shape[0]):
        outs.append(output[i].scatter(0, idx[i][:nums[i]].unsqueeze(-1).repeat(1, tgt.size(-1)), tgt[i][:nums[i]]))
    return torch.stack(outs)


@pytest.mark.parametrize("dtype,C", [(torch.float32, 512), (torch.bfloat16, 512), (torch.float32, 3)])
@pytest.mark.parametrize("with_nums", [False, True])
def test_gather_scatter_match_reference_expressions(dtype, C, with_nums):
    g = torch.Generator(device=DEV).manual_seed(5)
    B, S, K = 4, 300, 90
    x = torch.randn(B, S, C, device=DEV, generator=g).to(dtype).requires_grad_(True)
    score = torch.rand(B, S, device=DEV, generator=g)
    idx = torch.topk(score, K, dim=1)[1]                      # score order, as :218
    tgt = torch.randn(B, K, C, device=DEV, generator=g).to(dtype).requires_grad_(True)
    nums = torch.tensor([90, 17, 0, 64], device=DEV) if with_nums else None
    got_g = gather_tokens(x, idx)
    assert torch.equal(got_g, ref_gather(x, idx))
    got_s = scatter_tokens(x, idx, tgt, nums)
    want_s = ref_scatter(x, idx, tgt, nums.tolist() if with_nums else None)
    assert torch.equal(got_s, want_s)
    w = torch.randn_like(got_s)
    gx, gt = torch.autograd.grad((got_s.float() * w.float()).sum(), (x, tgt))
    rx, rt = torch.autograd.grad((want_s.float() * w.float()).sum(), (x, tgt))
    assert torch.equal(gx, rx) and torch.equal(gt, rt)
    w2 = torch.randn_like(got_g)
    (gx2,) = torch.autograd.grad((got_g.float() * w2.float()).sum(), (x,))
    (rx2,) = torch.autograd.grad((ref_gather(x, idx).float() * w2.float()).sum(), (x,))
    assert torch.equal(gx2, rx2)


def test_empty_and_cpu():
    x = torch.randn(2, 0, 8, device=DEV)
    idx = torch.zeros(2, 0, dtype=torch.long, device=DEV)
    assert gather_tokens(x, idx).shape == (2, 0, 8)
    with pytest.raises(RuntimeError, match="Not implemented on the CPU"):
        gather_tokens(torch.randn(1, 4, 8), torch.zeros(1, 2, dtype=torch.long))
