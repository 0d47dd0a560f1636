"""Token gather / scatter of the sparse encoders.

``gather_tokens(x, idx)`` is ``torch.gather(x, 1, idx.unsqueeze(-1).repeat(1, 1, C))``
(models/sparse/unimodal_sparse_deformable_transformer.py:425-431, multimodal twin :491-544)
without the repeated index; ``scatter_tokens(base, idx, src, nums)`` is the block at :443-448 --
``output.scatter(1, idx..., tgt)`` or, with per-item token counts, the Python loop of
``output[i].scatter(0, idx[i][:n_i]..., tgt[i][:n_i])`` -- as ONE launch for the batch with the
counts read on the device (no host sync, no per-item kernels).  Both are differentiable.
CUDA tensors only (no CPU path in this package)."""
import torch
from torch.autograd import Function
from torch.autograd.function import once_differentiable

from .. import _lib


def _stream(t):
    return torch.cuda.current_stream(t.device).cuda_stream


def _check(x, idx, nums):
    if not x.is_cuda:
        raise RuntimeError("Not implemented on the CPU")
    if idx.dtype != torch.int64 or idx.dim() != 2 or idx.shape[0] != x.shape[0]:
        raise RuntimeError("idx must be an int64 [B, K] tensor")
    if nums is not None and (nums.dtype != torch.int64 or nums.numel() != x.shape[0] or not nums.is_cuda):
        raise RuntimeError("nums must be an int64 CUDA tensor of B per-item counts")


def _gather(x, idx, nums):
    B, S = x.shape[:2]
    K = idx.shape[1]
    out = torch.empty((B, K) + tuple(x.shape[2:]), dtype=x.dtype, device=x.device)
    row = x[0, 0].numel() * x.element_size() if S > 0 else out[0, 0].numel() * out.element_size() if K > 0 else 1
    with torch.cuda.device_of(x):
        _lib.check(_lib.load().msda_token_gather(
            x.data_ptr(), idx.data_ptr(), nums.data_ptr() if nums is not None else None, out.data_ptr(),
            B, S, K, max(row, 1), _stream(x)))
    return out


def _scatter(base, idx, nums, src, zero_hit):
    B, S = base.shape[:2]
    K = idx.shape[1]
    out = torch.empty_like(base)
    inv = torch.empty(max(B * S, 1), dtype=torch.int32, device=base.device)
    row = base[0, 0].numel() * base.element_size() if S > 0 else 1
    with torch.cuda.device_of(base):
        _lib.check(_lib.load().msda_token_scatter(
            base.data_ptr(), idx.data_ptr(), nums.data_ptr() if nums is not None else None,
            src.data_ptr() if src is not None else None, out.data_ptr(), inv.data_ptr(),
            B, S, K, max(row, 1), 1 if zero_hit else 0, _stream(base)))
    return out


class _GatherTokens(Function):
    @staticmethod
    def forward(ctx, x, idx, nums):
        x, idx = x.contiguous(), idx.contiguous()
        _check(x, idx, nums)
        ctx.save_for_backward(idx, nums if nums is not None else idx.new_empty(0))
        ctx.has_nums, ctx.shape = nums is not None, x.shape
        return _gather(x, idx, nums)

    @staticmethod
    @once_differentiable
    def backward(ctx, g):
        idx, nums = ctx.saved_tensors
        zeros = torch.zeros(ctx.shape, dtype=g.dtype, device=g.device)
        return _scatter(zeros, idx, nums if ctx.has_nums else None, g.contiguous(), False), None, None


class _ScatterTokens(Function):
    @staticmethod
    def forward(ctx, base, idx, src, nums):
        base, idx, src = base.contiguous(), idx.contiguous(), src.contiguous()
        _check(base, idx, nums)
        if src.shape[:2] != idx.shape or src.shape[2:] != base.shape[2:] or src.dtype != base.dtype:
            raise RuntimeError("src must be [B, K, ...] with base's trailing shape and dtype")
        ctx.save_for_backward(idx, nums if nums is not None else idx.new_empty(0))
        ctx.has_nums = nums is not None
        return _scatter(base, idx, nums, src, False)

    @staticmethod
    @once_differentiable
    def backward(ctx, g):
        idx, nums = ctx.saved_tensors
        nums = nums if ctx.has_nums else None
        g = g.contiguous()
        return _scatter(g, idx, nums, None, True), None, _gather(g, idx, nums), None


def gather_tokens(x, idx, nums=None):
    """x [B, S, ...], idx [B, K] int64 -> [B, K, ...]; rows k >= nums[b] come back as zeros."""
    return _GatherTokens.apply(x, idx, nums)


def scatter_tokens(base, idx, src, nums=None):
    """A copy of base [B, S, ...] with rows idx[b, k] (k < nums[b]; all K when nums is None)
    replaced by src[b, k]; the indices of one batch item must be distinct (torch.topk's are)."""
    return _ScatterTokens.apply(base, idx, src, nums)
