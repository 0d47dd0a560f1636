#!/usr/bin/env python
"""CPU model of the CTA-tile plan (csrc/msda_tile.cuh): time-sort the queries as msda_sort_kernel
does, cut them into tiles of 64 and blocks of 4, and report the row-window statistics the kernels
will see.  Dev aid for capacity constants; needs no GPU.
   python tools/tile_stats.py [--workload ENC] [--dist model] [--sigma 0.5] [--N 2] [--order index|perm]"""
import argparse
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mfl_b200 import synthetic  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="ENC")
    ap.add_argument("--dist", default="model")
    ap.add_argument("--N", type=int, default=2)
    ap.add_argument("--tileq", type=int, default=64)
    ap.add_argument("--bins", type=int, default=4096)
    ap.add_argument("--cap", type=int, default=128)
    ap.add_argument("--blkcap", type=int, default=56)
    ap.add_argument("--box", type=int, default=8)
    ap.add_argument("--nosort", action="store_true")
    ap.add_argument("--hk", type=int, default=8)
    ap.add_argument("--sigma", type=float, default=None)
    args = ap.parse_args()
    kw = {}
    if args.sigma is not None:
        kw["sigma"] = args.sigma
    inp = synthetic.make_workload(args.workload, torch.float32, "cpu", seed=1234, dist=args.dist, N=args.N, **kw)
    loc = inp["loc"].numpy()          # [N, Lq, M, L, P, 2]
    shapes = inp["shapes"].numpy()
    lsi = inp["lsi"].numpy()
    N, Lq, M, L, P, _ = loc.shape
    W = shapes[:, 1]
    hk = min(M, args.hk)
    key = loc[:, :, :hk, 0, :, 0].mean(axis=(2, 3))
    bins = np.clip(np.floor(key * args.bins), 0, args.bins - 1).astype(np.int64)
    blk_rows, tile_rows, over_tile, over_blk, tiles = [], [], 0, 0, 0
    warp_rows, warp_rows_even, over_bwd = [], [], 0
    per_level = np.zeros(L)
    for b in range(N):
        perm = np.arange(Lq) if args.nosort else np.argsort(bins[b], kind="stable")
        for m in range(M):
            x = loc[b, perm, m, :, :, 0]                      # [Lq, L, P]
            w_im = x * W[None, :, None] - 0.5
            live = (w_im > -1) & (w_im < W[None, :, None])
            r0 = np.floor(w_im)
            lo = np.where(live, r0, 1e9).min(2)               # [Lq, L]
            hi = np.where(live, r0 + 1, -1e9).max(2)
            lo = np.maximum(lo, 0)
            hi = np.minimum(hi, W[None, :] - 1)
            for t0 in range(0, Lq, args.tileq):
                tiles += 1
                bad_bwd_warp = False
                tl, th = lo[t0:t0 + args.tileq], hi[t0:t0 + args.tileq]
                tlo, thi = tl.min(0), th.max(0)
                rows = np.where(thi >= tlo, thi - tlo + 1, 0)
                total = (np.ceil(rows / args.box) * args.box).sum()
                tile_rows.append(rows.sum())
                bad = total > args.cap
                # the kernels lay a warp's 4 blocks out with the warp's largest row count per level
                # (backward: rounded up to even): that sum is what must fit the block plan
                for w0 in range(0, len(tl), 16):
                    wr = np.zeros(L)
                    for q0 in range(w0, min(w0 + 16, len(tl)), 4):
                        bl, bh = tl[q0:q0 + 4].min(0), th[q0:q0 + 4].max(0)
                        wr = np.maximum(wr, np.where(bh >= bl, bh - bl + 1, 0))
                    warp_rows.append(wr.sum())
                    warp_rows_even.append((np.ceil(wr / 2) * 2).sum())
                    if (np.ceil(wr / 2) * 2).sum() > args.blkcap:
                        bad_bwd_warp = True
                for q0 in range(0, len(tl), 4):
                    bl, bh = tl[q0:q0 + 4].min(0), th[q0:q0 + 4].max(0)
                    br = np.where(bh >= bl, bh - bl + 1, 0)
                    blk_rows.append(br.sum())
                    per_level += br
                    if br.sum() > args.blkcap:
                        bad = True
                        over_blk += 1
                if bad_bwd_warp or total > args.cap:
                    over_bwd += 1
                if bad:
                    over_tile += 1
                    if over_tile <= 6:
                        print("overflow tile b=%d m=%d q0=%d rows/level=%s key range %.4f..%.4f" % (
                            b, m, t0, rows.astype(int).tolist(), key[b, perm[t0]], key[b, perm[min(t0 + args.tileq, Lq) - 1]]))
    blk_rows, tile_rows = np.array(blk_rows), np.array(tile_rows)
    print("tiles %d, overflow %d (%.1f%%), blocks over cap %d" % (tiles, over_tile, 100.0 * over_tile / tiles, over_blk))
    print("rows per block: mean %.1f p50 %d p90 %d p99 %d max %d; per level %s" % (
        blk_rows.mean(), np.percentile(blk_rows, 50), np.percentile(blk_rows, 90), np.percentile(blk_rows, 99),
        blk_rows.max(), (per_level / len(blk_rows)).round(1).tolist()))
    wr, we = np.array(warp_rows), np.array(warp_rows_even)
    print("plan rows per warp (sum over levels of the largest block): fwd mean %.1f p99 %d max %d; bwd (even) mean %.1f "
          "p99 %d max %d; tiles the backward sends to the per-tap body: %d (%.1f%%)" % (
              wr.mean(), np.percentile(wr, 99), wr.max(), we.mean(), np.percentile(we, 99), we.max(), over_bwd,
              100.0 * over_bwd / tiles))
    print("rows per tile: mean %.1f p90 %d p99 %d max %d" % (tile_rows.mean(), np.percentile(tile_rows, 90),
                                                              np.percentile(tile_rows, 99), tile_rows.max()))


if __name__ == "__main__":
    main()
