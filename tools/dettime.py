#!/usr/bin/env python
"""Times the deterministic backward (MSDA_FLAG_DETERMINISTIC) against the default one at ENC, through
the Python shim (allocation + zero-fill inside), CUDA events, median of 20.  Dev aid."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mfl_b200  # noqa: E402,F401
from mfl_b200 import MultiScaleDeformableAttention as MSDA  # noqa: E402
from mfl_b200 import _lib, synthetic  # noqa: E402

dev = "cuda"
for dtype in (torch.float32, torch.bfloat16):
    sets = [synthetic.make_workload("ENC", dtype, dev, seed=5 + r, dist="model", value="normal") for r in range(3)]
    for name, fl in (("default", 0), ("deterministic", _lib.MSDA_FLAG_DETERMINISTIC),
                     ("deterministic, row windows", _lib.MSDA_FLAG_DETERMINISTIC | _lib.MSDA_FLAG_NO_TILE),
                     ("deterministic, per-tap", _lib.MSDA_FLAG_DETERMINISTIC | _lib.MSDA_FLAG_NO_WINDOW)):
        ts = []
        for i in range(25):
            c = sets[i % 3]
            a = [c[k] for k in ("value", "shapes", "lsi", "loc", "aw", "grad_out")]
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            MSDA.ms_deform_attn_backward(*a, 64, flags=fl)
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1) * 1e3)
        print("%-9s %-28s backward %.1f us" % (str(dtype)[6:], name, sorted(ts[5:])[10]))
