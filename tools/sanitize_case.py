#!/usr/bin/env python
"""Small launches of every kernel family, for compute-sanitizer:

  PYTORCH_NO_CUDA_MEMORY_CACHING=1 compute-sanitizer --tool memcheck  python tools/sanitize_case.py
  PYTORCH_NO_CUDA_MEMORY_CACHING=1 compute-sanitizer --tool racecheck python tools/sanitize_case.py

Covers the window kernels (forced; blocks that fit, blocks that fall back, ragged query
counts, fewer than 4 levels), the per-tap kernels and the generic ones, fp32 and bf16,
both paddings.  Prints one line per launch pair; the sanitizer's own summary is the result.
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import mfl_b200  # noqa: E402,F401
from mfl_b200 import MultiScaleDeformableAttention as MSDA  # noqa: E402
from mfl_b200 import _lib, synthetic  # noqa: E402

CASES = [
    # N, T, M, D, Lq (None = encoder tokens), dist
    (1, [48, 24, 12, 6], 2, 64, None, "model"),     # Lq = 90: ragged unit, windows fit
    (1, [48, 24, 12, 6], 2, 64, None, "uniform"),   # every block falls back
    (1, [30, 15], 2, 32, None, "model"),            # two levels, D = 32
    (1, [20, 10, 5, 3], 2, 64, 3, "model"),         # fewer queries than one block
    (1, [40, 20, 10, 5], 1, 64, 21, "spill"),       # taps off both ends
]
FLAGS = [("default", 0), ("generic", _lib.MSDA_FLAG_FORCE_GENERIC), ("per-tap", _lib.MSDA_FLAG_NO_WINDOW),
         ("window", _lib.MSDA_FLAG_FORCE_WINDOW)]


def main():
    dev = "cuda"
    for ci, (N, T, M, D, Lq, dist) in enumerate(CASES):
        enc = Lq is None
        for dtype in (torch.float32, torch.bfloat16):
            inp = synthetic.make_inputs(N, sum(T) if enc else Lq, M, D, 4, [(1, t) for t in T], dtype, dev,
                                        seed=3 + ci, dist=dist, value="normal", enc_ref=enc)
            a = [inp[k] for k in ("value", "shapes", "lsi", "loc", "aw")]
            for padding in ("zeros", "border"):
                ref = None
                for name, fl in FLAGS:
                    out = MSDA.ms_deform_attn_forward(*a, 64, padding_mode=padding, flags=fl)
                    gv, gl, ga = MSDA.ms_deform_attn_backward(*a, inp["grad_out"], 64, padding_mode=padding,
                                                              flags=fl)
                    torch.cuda.synchronize()
                    cur = [t.float() for t in (out, gv, ga)]
                    if ref is None:
                        ref = cur
                    err = max(float((x - y).abs().max() / y.abs().max().clamp_min(1e-30)) for x, y in zip(cur, ref))
                    print("case %d %-8s %-6s %-8s max rel diff vs default %.2e" % (
                        ci, str(dtype).split(".")[-1], padding, name, err), flush=True)
    print("done")


if __name__ == "__main__":
    main()
