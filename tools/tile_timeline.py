#!/usr/bin/env python
"""Phase timeline of the forward tile kernel (or, with --bwd [--flags 128], of the backward) from a
-DMSDA_TILE_TIMING build (variants/timing.so, or --lib): globaltimer stamps of thread 0 of every CTA.
Dev aid."""
import ctypes
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mfl_b200 import _lib  # noqa: E402

import argparse
ap = argparse.ArgumentParser()
ap.add_argument("--bwd", action="store_true")
ap.add_argument("--flags", type=int, default=0)
ap.add_argument("--lib", default=os.path.join(ROOT, "variants", "timing.so"))
args = ap.parse_args()
_lib.LIB_PATH = os.path.abspath(args.lib)
import bench  # noqa: E402
from mfl_b200 import synthetic  # noqa: E402

dev = torch.device("cuda", 0)
ring = [synthetic.make_workload("ENC", torch.float32, dev, seed=1234 + r, dist="model", value="normal") for r in range(4)]
bufs = [bench.OpBuffers(i, "zeros", 0) for i in ring]
st = torch.cuda.current_stream().cuda_stream
for b_ in bufs:
    b_.set_flags(args.flags)
for i in range(9):
    bufs[i % 4].fwd(st)
    if args.bwd:
        bufs[i % 4].zero()
        bufs[i % 4].bwd(st)
torch.cuda.synchronize()
lib = _lib.load()
n = 3072
host = (ctypes.c_ulonglong * (8 * n))()
lib.msda_debug_tile_times.argtypes = [ctypes.c_void_p, ctypes.c_int]
assert lib.msda_debug_tile_times(host, 8 * n) == 0
t = np.array(host, dtype=np.float64).reshape(n, 8)
t0 = t[:, 0].min()
if args.bwd:
    print("BACKWARD kernel, flags", args.flags, "(phases: setup | grad_out requested, points+geometry or saved header | "
          "barrier / issue | scatter or nothing | wait for boxes | row loops | epilogue)")
names = ["start->setup(perm)", "setup->geometry done", "geometry->barrier", "barrier->plan built(TMA issued)",
         "plan->V landed", "V->rows done", "rows->stored"]
d = np.diff(t, axis=1)
print("kernel span %.1f us, CTA life mean %.2f us" % ((t[:, 7].max() - t0) / 1e3, (t[:, 7] - t[:, 0]).mean() / 1e3))
for k, nm in enumerate(names):
    print("  %-34s mean %6.0f ns  p10 %6.0f  p90 %6.0f" % (nm, d[:, k].mean(), np.percentile(d[:, k], 10), np.percentile(d[:, k], 90)))
start = np.sort(t[:, 0] - t0)
print("CTA start times (us): first wave until %.1f, median %.1f, last %.1f" % (start[591] / 1e3, np.median(start) / 1e3, start[-1] / 1e3))

hs = (ctypes.c_ulonglong * (8 * 16))()
lib.msda_debug_sort_times.argtypes = [ctypes.c_void_p, ctypes.c_int]
assert lib.msda_debug_sort_times(hs, 8 * 16) == 0
ts = np.array(hs, dtype=np.float64).reshape(16, 8)[:, :5]
print("sort kernel phases (ns), mean over 16 CTAs: zero %s | keys+hist %s | scan %s | scatter %s ; first start -> last end %.1f us; sort end -> first tile CTA start %.1f us" % (
    *[("%.0f" % x) for x in np.diff(ts, axis=1).mean(0)], (ts[:, 4].max() - ts[:, 0].min()) / 1e3, (t[:, 0].min() - ts[:, 4].max()) / 1e3))
