#!/usr/bin/env python
"""Time msda_frontend_forward / _backward through the C ABI at the ENC and DEC module shapes.
   python tools/frontbench.py [--iters 50] [--dtype f32|bf16]"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mfl_b200  # noqa: E402,F401
from mfl_b200 import _lib, synthetic  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=50)
    ap.add_argument("--dtype", default="f32")
    ap.add_argument("--ring", type=int, default=6)
    args = ap.parse_args()
    dtype = {"f32": torch.float32, "bf16": torch.bfloat16}[args.dtype]
    dt = {"f32": 0, "bf16": 2}[args.dtype]
    peak = 6550.0
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        pass
    lib = _lib.load()
    st = torch.cuda.current_stream().cuda_stream
    for wl in ("ENC", "DEC"):
        w = synthetic.WORKLOADS[wl]
        N, M, L, P = w["N"], w["M"], len(w["T"]), w["P"]
        Lq = w["Lq"] or sum(w["T"])
        sizes = torch.tensor(w["T"], dtype=torch.long, device="cuda")
        esz = torch.empty((), dtype=dtype).element_size()
        n = N * Lq * M * L * P
        for rd, need_ref in ((1, False), (2, True)):
            sets = []
            for r in range(args.ring):   # rotate over > L2 worth of buffers
                g = torch.Generator(device="cuda").manual_seed(r)
                mk = lambda *s: torch.randn(*s, device="cuda", generator=g).to(dtype)
                sets.append(dict(off=mk(N, Lq, M, L, P), lg=mk(N, Lq, M, L * P), ref=torch.rand(N, Lq, L, rd, device="cuda", generator=g).to(dtype),
                                 loc=torch.empty(N, Lq, M, L, P, 2, device="cuda", dtype=dtype), aw=torch.empty(N, Lq, M, L, P, device="cuda", dtype=dtype),
                                 gloc=mk(N, Lq, M, L, P, 2), gaw=mk(N, Lq, M, L, P), goff=torch.empty(N, Lq, M, L, P, device="cuda", dtype=dtype),
                                 glg=torch.empty(N, Lq, M, L, P, device="cuda", dtype=dtype), gref=torch.empty(N, Lq, L, rd, device="cuda", dtype=dtype)))

            def fwd(s):
                _lib.check(lib.msda_frontend_forward(s["off"].data_ptr(), s["lg"].data_ptr(), s["ref"].data_ptr(), sizes.data_ptr(),
                                                     s["loc"].data_ptr(), s["aw"].data_ptr(), N, Lq, M, L, P, rd, dt, st))

            def bwd(s):
                _lib.check(lib.msda_frontend_backward(s["off"].data_ptr(), s["aw"].data_ptr(), s["ref"].data_ptr(), sizes.data_ptr(),
                                                      s["gloc"].data_ptr(), s["gaw"].data_ptr(), s["goff"].data_ptr(), s["glg"].data_ptr(),
                                                      s["gref"].data_ptr() if need_ref else None, N, Lq, M, L, P, rd, dt, st))
            for i in range(6):
                fwd(sets[i % len(sets)]); bwd(sets[i % len(sets)])
            ev = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(args.iters)]
            torch.cuda.synchronize()
            for i in range(args.iters):
                s = sets[i % len(sets)]
                ev[i][0].record(); fwd(s); ev[i][1].record(); bwd(s); ev[i][2].record()
            torch.cuda.synchronize()
            tf, tb = (sorted(e[k].elapsed_time(e[k + 1]) for e in ev)[len(ev) // 2] * 1e3 for k in range(2))
            bf, bb = esz * n * 5, esz * n * 7          # fwd: off, lg -> loc(2), aw ; bwd: off, aw, gloc(2), gaw -> goff, glg
            print("%s %s ref_dim %d  fwd %6.1f us %5.0f GB/s (%.2f of peak)   bwd %6.1f us %5.0f GB/s (%.2f)" % (
                wl, args.dtype, rd, tf, bf / tf / 1e3, bf / tf / 1e3 / peak, tb, bb / tb / 1e3, bb / tb / 1e3 / peak), flush=True)


if __name__ == "__main__":
    main()
