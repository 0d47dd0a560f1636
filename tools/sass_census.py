#!/usr/bin/env python
"""Opcode census of the tile / sort / plan / apply kernels of libmsda_b200.so (cuobjdump -sass,
sm_100a): the evidence that the value tiles move by TMA (UTMALDG = cp.async.bulk.tensor, UBLKCP =
cp.async.bulk, UTMAPF = prefetch, SYNCS = mbarrier operations), that grad_value goes out as
red.global.add.v4.f32 (REDG) and that the FMAs are the packed FFMA2.  Needs no GPU.
   python tools/sass_census.py [path/to/libmsda_b200.so] > profiles/<round>_sass_tile.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "multimodal-feature-learning_b200", "libmsda_b200.so")
KEEP = ("UTMALDG", "UTMASTG", "UTMAREDG", "UTMAPF", "UBLKCP", "SYNCS", "LDGSTS", "LDS", "STS", "LDG", "STG", "REDG",
        "RED", "ATOMG", "ATOMS", "FFMA2", "FFMA", "SHFL", "BAR")
WANT = re.compile(r"msda_(fwd|bwd)_tile_kernel|msda_sort|msda_plan_kernel|msda_fwd_apply_kernel")

sass = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True, check=True).stdout
kernels, name = collections.OrderedDict(), None
for line in sass.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        name = m.group(1)
        kernels[name] = collections.Counter()
        continue
    m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)", line)
    if m and name:
        kernels[name]["_all"] += 1
        if m.group(1) in KEEP:
            kernels[name][m.group(1)] += 1

print("SASS opcode census of %s (cuobjdump -sass, sm_100a), kernels matching %s." % (os.path.basename(so), WANT.pattern))
print("UTMALDG = cp.async.bulk.tensor, UBLKCP = cp.async.bulk, UTMAPF = cp.async.bulk.prefetch.tensor, SYNCS = mbarrier ops,")
print("REDG = red.global.add (v4.f32 in the tile backward), FFMA2 = packed fp32 FMA.  Static counts (loops are not unrolled")
print("where the count is small: the TMA box-issue loop is one UTMALDG in a loop since the last build).\n")
total = collections.Counter()
for k, c in kernels.items():
    for op, n in c.items():
        total[op] += n
    if WANT.search(k):
        print(k[:150])
        print("    %d instructions  %s" % (c["_all"], {op: c[op] for op in KEEP if c[op]}))
print("\nwhole library (%d kernels): %s" % (len(kernels), {op: total[op] for op in KEEP if total[op]}))
