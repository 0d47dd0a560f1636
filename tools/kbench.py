#!/usr/bin/env python
"""Developer benchmark: per-kernel CUDA-event timings of the C-ABI hot path for the named
workloads, with and without the row-window kernels, plus a cross-check of the two paths.

  python tools/kbench.py [--workloads ENC,DEC] [--dists model,uniform] [--dtypes f32,bf16] [--iters 50]
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

import bench  # noqa: E402
from mfl_b200 import _lib, synthetic  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workloads", default="ENC,DEC")
    ap.add_argument("--dists", default="model,uniform")
    ap.add_argument("--dtypes", default="f32,bf16")
    ap.add_argument("--iters", type=int, default=50)
    ap.add_argument("--flags", default="0,32,4", help="0 = default (tile), 32 = no tile (row windows), 4 = per-tap")
    ap.add_argument("--N", type=int, default=0, help="override the workload's batch size")
    ap.add_argument("--ring", type=int, default=4, help="distinct input sets rotated (1 = L2-resident inputs)")
    ap.add_argument("--lib", default="", help="A/B aid: load this build of libmsda_b200.so instead of the in-tree one")
    args = ap.parse_args()
    if args.lib:
        from mfl_b200 import _lib as _l
        _l.LIB_PATH = os.path.abspath(args.lib)
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    peak = 6550.0
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        pass
    st = torch.cuda.current_stream().cuda_stream
    for wl in args.workloads.split(","):
        for dist in args.dists.split(","):
            for dts in args.dtypes.split(","):
                dtype = {"f32": torch.float32, "bf16": torch.bfloat16}[dts]
                ring = [synthetic.make_workload(wl, dtype, dev, seed=1234 + r, dist=dist, value="normal", N=args.N or None)
                        for r in range(args.ring)]
                outs = {}
                for fl in [int(x) for x in args.flags.split(",")]:
                    bufs = [bench.OpBuffers(i, "zeros", 0) for i in ring]
                    for b in bufs:
                        b.set_flags(b.flags | fl)
                    b0 = bufs[0]
                    fb, bb = synthetic.algorithmic_bytes(b0.N, b0.S, b0.Lq, b0.M, b0.D, b0.L, b0.P,
                                                         ring[0]["value"].element_size())
                    for i in range(8):
                        b = bufs[i % len(bufs)]
                        b.fwd(st); b.zero(); b.bwd(st)
                    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(4)] for _ in range(args.iters)]
                    torch.cuda.synchronize()
                    for i in range(args.iters):
                        b = bufs[i % len(bufs)]
                        ev[i][0].record(); b.fwd(st)
                        ev[i][1].record(); b.zero()
                        ev[i][2].record(); b.bwd(st)
                        ev[i][3].record()
                    torch.cuda.synchronize()
                    seg = [sorted(e[k].elapsed_time(e[k + 1]) for e in ev)[len(ev) // 2] * 1e3 for k in range(3)]
                    print("%-4s %-8s %-5s flags=%d  fwd %7.1f us (%.3f of peak)  zero %6.1f us  bwd %7.1f us (%.3f)  "
                          "step %.3f" % (wl, dist, dts, fl, seg[0], fb / seg[0] / 1e3 / peak, seg[1], seg[2],
                                         bb / seg[2] / 1e3 / peak, (fb + bb) / sum(seg) / 1e3 / peak), flush=True)
                    outs[fl] = [t.float().clone() for t in (b0.out, b0.ws if b0.ws_bytes else b0.gv, b0.gl, b0.ga)]
                    # b0 ran last with i % 4 == 0 only if iters % 4 == 1; rerun to be sure
                    b0.fwd(st); b0.zero(); b0.bwd(st)
                    torch.cuda.synchronize()
                    outs[fl] = [t.float().clone() for t in (b0.out, b0.ws if b0.ws_bytes else b0.gv, b0.gl, b0.ga)]
                ks = list(outs)
                if len(ks) > 1:
                    errs = [float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))
                            for a, b in zip(outs[ks[0]], outs[ks[1]])]
                    print("     cross-check flags %d vs %d: rel err out/gvalue/gloc/gaw = %s" % (
                        ks[0], ks[1], " ".join("%.2e" % e for e in errs)), flush=True)


if __name__ == "__main__":
    main()
