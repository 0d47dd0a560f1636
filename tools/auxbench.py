#!/usr/bin/env python
"""Developer benchmark of the kernels around the op: decoder attention map
scatter, sparse-encoder token gather / scatter, encoder input preparation -- each against the chain of
torch kernels the reference runs for it (re-expressed here with stock torch ops), CUDA events, inputs
rotated over a ring larger than L2 where they are large enough to matter.  Timed through the Python
API: at these sizes every line is mostly host enqueue time (one launch vs a chain of launches).

  python tools/auxbench.py [--iters 50]
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

import mfl_b200  # noqa: E402,F401
from mfl_b200 import utils  # noqa: E402

DEV = "cuda"


def timed(fn, ring, iters):
    for a in ring:
        fn(*a)
    torch.cuda.synchronize()
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(2)] for _ in range(iters)]
    for i in range(iters):
        ev[i][0].record()
        fn(*ring[i % len(ring)])
        ev[i][1].record()
    torch.cuda.synchronize()
    return sorted(a.elapsed_time(b) for a, b in ev)[iters // 2] * 1e3


def torch_dam(T, lsi, loc, aw):
    """The shape of utils/dam.py:20-73 with stock ops (two taps per point, index_add into the flat grid):
    a timing stand-in, not a parity reference (tests/test_module_dam.py has the golden from the reference)."""
    B, NL, Q, M, L, P, _ = loc.shape
    S = int(T.sum())
    x = loc[..., 0] * T.view(1, 1, 1, 1, L, 1).float() - 0.5
    lo = x.floor()
    frac = x - lo
    out = torch.zeros(B * NL * M, S, device=loc.device)
    row = torch.arange(B * NL, device=loc.device).view(B, NL, 1, 1, 1, 1) * M + torch.arange(M, device=loc.device).view(1, 1, 1, M, 1, 1)
    for off, w in ((0, aw * (1 - frac)), (1, aw * frac)):
        col = lo.long() + off
        ok = (col >= 0) & (col < T.view(1, 1, 1, 1, L, 1))
        idx = (row * S + lsi.view(1, 1, 1, 1, L, 1) + col.clamp(min=0)).expand_as(w)
        out.view(-1).index_add_(0, idx[ok], w[ok])
    return out.view(B, NL, M, S)


def torch_gather(x, idx):
    return torch.gather(x, 1, idx.unsqueeze(-1).repeat(1, 1, x.size(-1)))


def torch_scatter(base, idx, src, nums):
    out = []
    for i in range(base.shape[0]):
        n = int(nums[i])
        out.append(base[i].scatter(0, idx[i][:n].unsqueeze(-1).repeat(1, base.size(-1)), src[i][:n]))
    return torch.stack(out)


def torch_prepare(srcs, masks, poss, level_embed):
    fs, fp, fm = [], [], []
    for lvl, (s, m, p) in enumerate(zip(srcs, masks, poss)):
        fs.append(s.transpose(1, 2))
        fp.append(p.transpose(1, 2) + level_embed[lvl].view(1, 1, -1))
        fm.append(m)
    vr = torch.stack([(~m).sum(1).float() / m.shape[1] for m in masks], 1)
    return torch.cat(fs, 1), torch.cat(fp, 1), torch.cat(fm, 1), vr


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=50)
    args = ap.parse_args()
    g = torch.Generator(device=DEV).manual_seed(7)
    T_list = [800, 400, 200, 100]
    T = torch.tensor(T_list, device=DEV)
    lsi = torch.cat((T.new_zeros(1), T.cumsum(0)[:-1]))

    # decoder attention map: 16 videos x 6 layers x 100 queries x 8 heads x 4 levels x 4 points
    ring = []
    for _ in range(4):
        loc = torch.rand(16, 6, 100, 8, 4, 4, 1, device=DEV, generator=g)
        aw = torch.softmax(torch.randn(16, 6, 100, 8, 16, device=DEV, generator=g), -1).view(16, 6, 100, 8, 4, 4)
        ring.append((T, lsi, loc, aw))
    a = timed(utils.attn_map_to_flat_grid, ring, args.iters)
    b = timed(torch_dam, ring, args.iters)
    print("attention map (1.23 M points -> [16, 6, 8, 1500]):  kernel %7.1f us   torch ops %8.1f us" % (a, b))

    # sparse encoder: keep 30 % of 1500 tokens, C = 512
    B, S, C, K = 16, 1500, 512, 450
    ring_g, ring_s = [], []
    for _ in range(6):
        x = torch.randn(B, S, C, device=DEV, generator=g)
        idx = torch.stack([torch.randperm(S, device=DEV, generator=g)[:K] for _ in range(B)])
        src = torch.randn(B, K, C, device=DEV, generator=g)
        nums = torch.full((B,), K, device=DEV, dtype=torch.long)
        ring_g.append((x, idx))
        ring_s.append((x, idx, src, nums))
    a = timed(lambda x, i: utils.gather_tokens(x, i), ring_g, args.iters)
    b = timed(torch_gather, ring_g, args.iters)
    print("token gather  (16 x 450 of 1500 tokens x 512, %.1f MB moved):   kernel %7.1f us   torch.gather %8.1f us" % (2 * B * K * C * 4 / 1e6, a, b))
    a = timed(lambda x, i, s_, n: utils.scatter_tokens(x, i, s_, n), ring_s, args.iters)
    b = timed(torch_scatter, ring_s, args.iters)
    print("token scatter (same, %.1f MB moved):                          kernel %7.1f us   per-item torch loop %8.1f us" % ((2 * B * S * C + B * K * C) * 4 / 1e6, a, b))

    # encoder inputs: 4 levels of [16, 512, T_l] + masks + position embeddings
    ring_p = []
    le = torch.randn(4, 512, device=DEV, generator=g)
    for _ in range(4):
        srcs = [torch.randn(16, 512, t, device=DEV, generator=g) for t in T_list]
        poss = [torch.randn(16, 512, t, device=DEV, generator=g) for t in T_list]
        masks = [torch.rand(16, t, device=DEV, generator=g) > 0.9 for t in T_list]
        ring_p.append((srcs, masks, poss, le))
    a = timed(utils.prepare_encoder_inputs, ring_p, args.iters)
    b = timed(torch_prepare, ring_p, args.iters)
    print("prepare_encoder_inputs (2 x 49 MB in, 2 x 49 MB out):          kernels %6.1f us   torch ops %8.1f us" % (a, b))


if __name__ == "__main__":
    main()
