#!/usr/bin/env python
"""Rewrite the table of DESIGN.md section 6 (and the sentence under it) from a bench line:
   python profiles/design_table.py profiles/r2c_bench_enc_f32_1gpu.json"""
import json
import re
import sys

d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
s = d["secondary"]


def row(k):
    v = s[k]
    return v["fwd_ms"] * 1e3, v["bwd_ms"] * 1e3, v["step_ms"] * 1e3, v["step_frac"]


def fmt(label, fam, k):
    return "| %s | %s | %.1f | %.1f | %.1f | %.3f |" % ((label, fam) + row(k))


def fmtm(label, fam, ks):
    r = [row(k) for k in ks]
    return "| %s | %s | %s | %s | %s | %s |" % (
        label, fam, " / ".join("%.1f" % x[0] for x in r), " / ".join("%.1f" % x[1] for x in r),
        " / ".join("%.0f" % x[2] for x in r), " / ".join("%.3f" % x[3] for x in r))


r, rep = d["roofline"], d["repeats"]
L = ["| **ENC f32 (headline)** | tile | %.1f | %.1f | **%.1f** (min %.1f, max %.1f) | **%.3f** (bwd kernel %.3f) |" % (
    r["fwd_kernel"]["kernel_ms"] * 1e3, r["kernel_ms"] * 1e3, d["ms_per_step"] * 1e3, rep["ms_per_step_min"] * 1e3,
    rep["ms_per_step_max"] * 1e3, r["step"]["frac"], r["frac"])]
L.append(fmt("ENC f32, zero-fill serial instead of overlapped", "tile", "ENC_f32_zero_fill_serial"))
L.append(fmt("ENC f32, row windows (`MSDA_FLAG_NO_TILE`, round 1's default)", "window", "ENC_f32_no_tile(row windows, round 1 default)"))
L.append(fmt("ENC f32, plan + apply split (`MSDA_FLAG_PLAN`; backward = tile)", "tile", "ENC_f32_plan_apply_split"))
L.append(fmt("ENC f32, per-tap kernels (`MSDA_FLAG_NO_WINDOW`)", "fast", "ENC_f32_per_tap"))
L.append(fmt("ENC f32, inputs L2-resident (1 set)", "tile", "ENC_f32_l2warm"))
L.append(fmt("**ENC f32, permuted query list** (sparse encoders)", "tile", "ENC_f32_permuted"))
L.append(fmt("same, row windows", "window", "ENC_f32_permuted_no_tile"))
L.append(fmt("same, per-tap", "fast", "ENC_f32_permuted_per_tap"))
L.append(fmt("ENC f32, offsets + N(0, 2²) rows (σ = 2)", "tile", "ENC_f32_sigma2"))
L.append(fmtm("ENC f32, σ = 8 / σ = 32 / uniform locations (every tile falls back)", "tile",
              ["ENC_f32_sigma8", "ENC_f32_sigma32", "ENC_f32_uniform"]))
L.append(fmt("ENC bf16", "tile", "ENC_bf16"))
L.append(fmtm("DEC f32 (N=64, Lq=100) / DEC bf16", "fast", ["DEC_f32", "DEC_bf16"]))
mm = ["MM_VV_f32", "MM_AA_f32", "MM_AV_f32", "MM_VA_f32"]
L.append(fmtm("MM_VV / MM_AA / MM_AV / MM_VA f32 (the four calls of a multimodal layer)", "/".join(s[k]["family"] for k in mm), mm))
L.append("| configs[0] shape on the GPU (N=2, Lq=100, T=[300,150,75,38]) | fast | %.1f | %.1f | %.1f | %.3f (launch-bound) |"
         % row("CPU_shape_f32(configs[0] on the GPU)"))
k = s["kernel_to_beat"]
tot = k["fwd_ms"] + k["bwd_ms"]
L.append("| kernel to beat: the reference's CUDA op recompiled for sm_100a (`oracle/_ref`) | – | %.0f | %.0f | %.0f | %.3f |" % (
    k["fwd_ms"] * 1e3, k["bwd_ms"] * 1e3, tot * 1e3, 356.352e6 / (tot * 1e-3) / 1e9 / r["peak"]))
t = open("DESIGN.md").read()
a, b = t.index("| **ENC f32 (headline)** |"), t.index("\nAgainst round 1 (175.4")
t = t[:a] + "\n".join(L) + "\n" + t[b:]
t = re.sub(r"Against round 1 \(175\.4 µs, 0\.31\): \*\*[0-9.]+ µs, [0-9.]+\*\* \(backward kernel [0-9.]+ µs = [0-9.]+, forward kernel \+ sort\n[0-9.]+ µs\)",
           "Against round 1 (175.4 µs, 0.31): **%.1f µs, %.3f** (backward kernel %.1f µs = %.3f, forward kernel + sort\n%.1f µs)" % (
               d["ms_per_step"] * 1e3, r["step"]["frac"], r["kernel_ms"] * 1e3, r["frac"], r["fwd_kernel"]["kernel_ms"] * 1e3), t)
open("DESIGN.md", "w").write(t)
print("step %.1f us, %.3f; module call %s" % (d["ms_per_step"] * 1e3, r["step"]["frac"], s["ENC_f32_module_call"]))
