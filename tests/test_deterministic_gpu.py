"""MSDA_FLAG_DETERMINISTIC: grad_value accumulated as 64-bit fixed point with integer atomics
(csrc/msda_det.cuh).  Checks: (1) the same tolerances against the C oracle as every other path,
(2) bitwise identical results over repeated runs on a contended problem (many taps per row),
(3) torch.use_deterministic_algorithms(True) selects it through the autograd Function,
(4) extreme magnitudes: the scale follows max|grad_out| * max|attn_weight|."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

import mfl_b200  # noqa: E402,F401
from mfl_b200 import MultiScaleDeformableAttention as MSDA  # noqa: E402
from mfl_b200 import _lib, synthetic  # noqa: E402
from mfl_b200.functions.ms_deform_attn_func import ms_deform_attn  # noqa: E402

DEV = "cuda"
DET = _lib.MSDA_FLAG_DETERMINISTIC


def rel_err(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30)) if b.size else 0.0


def bwd(inp, padding, flags):
    a = [inp[k] for k in ("value", "shapes", "lsi", "loc", "aw", "grad_out")]
    gv, gl, ga = MSDA.ms_deform_attn_backward(*a, 64, padding_mode=padding, flags=flags)
    torch.cuda.synchronize()
    return gv, gl, ga


CASES = [
    (2, 37, 8, 64, 4, [(1, 38), (1, 19), (1, 10), (1, 5)]),
    (2, 50, 4, 32, 4, [(7, 9), (4, 5), (2, 3), (1, 2)]),
    (1, 13, 2, 30, 3, [(5, 5), (3, 3)]),
]


@pytest.mark.parametrize("case", CASES, ids=["temporal", "2d", "oddD"])
@pytest.mark.parametrize("padding", ["zeros", "border"])
@pytest.mark.parametrize("dtype", [torch.float32, torch.float64, torch.bfloat16], ids=["f32", "f64", "bf16"])
def test_deterministic_matches_oracle(oracle_c, case, padding, dtype):
    N, Lq, M, D, P, shapes = case
    inp = synthetic.make_inputs(N, Lq, M, D, P, shapes, dtype, DEV, seed=5, dist="spill", value="normal")
    f = {k: (v.double() if v.is_floating_point() else v) for k, v in inp.items()}
    if dtype != torch.float64:
        f = {k: (v.float() if v.is_floating_point() else v) for k, v in inp.items()}
    want = oracle_c.backward(*[f[k] for k in ("value", "shapes", "lsi", "loc", "aw")], f["grad_out"], padding)
    got = [t.double().cpu().numpy() for t in bwd(inp, padding, DET)]
    tol = {torch.float32: 1e-4, torch.float64: 1e-9, torch.bfloat16: 2e-2}[dtype]
    for g, w, nm in zip(got, want, ("grad_value", "grad_loc", "grad_aw")):
        assert rel_err(g, w) <= tol, nm


def test_bitwise_reproducible_under_contention():
    """2400 queries x 8 points x 2 taps onto 24 + 12 rows: ~2000 addends per grad_value element."""
    inp = synthetic.make_inputs(2, 1200, 4, 64, 4, [(1, 24), (1, 12)], torch.float32, DEV, seed=6, dist="uniform",
                                value="normal")
    runs = [bwd(inp, "zeros", DET) for _ in range(4)]
    for r in runs[1:]:
        for a, b in zip(runs[0], r):
            assert torch.equal(a, b)
    # the default path agrees within rounding; its grad_value is the non-reproducible one
    dflt = bwd(inp, "zeros", 0)
    assert rel_err(runs[0][0].cpu(), dflt[0].cpu()) <= 1e-5
    assert torch.equal(runs[0][1], dflt[1]) or rel_err(runs[0][1].cpu(), dflt[1].cpu()) <= 1e-5


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16], ids=["f32", "bf16"])
def test_window_and_per_tap_families_are_each_reproducible(oracle_c, dtype):
    """An encoder-like call on the CTA-tile kernels, on the row-window kernels (with blocks that
    fall back) and on the per-tap kernels: all deterministic, all within the oracle tolerance."""
    T = [96, 48, 24, 12]
    inp = synthetic.make_inputs(2, sum(T), 8, 64, 4, [(1, t) for t in T], dtype, DEV, seed=9, dist="model",
                                value="normal", enc_ref=True)
    f = {k: (v.float() if v.is_floating_point() else v) for k, v in inp.items()}
    want = oracle_c.backward(*[f[k] for k in ("value", "shapes", "lsi", "loc", "aw")], f["grad_out"], "zeros")
    lib = _lib.load()
    for fl, fam in ((_lib.MSDA_FLAG_FORCE_WINDOW, "tile"),
                    (_lib.MSDA_FLAG_FORCE_WINDOW | _lib.MSDA_FLAG_NO_TILE, "window"),
                    (_lib.MSDA_FLAG_NO_WINDOW, "fast")):
        assert lib.msda_kernel_choice(2, sum(T), 8, 64, 4, sum(T), 4, 0 if dtype == torch.float32 else 2, fl).decode() == fam
        a = bwd(inp, "zeros", DET | fl)
        b = bwd(inp, "zeros", DET | fl)
        assert all(torch.equal(x, y) for x, y in zip(a, b)), fam
        assert rel_err(a[0].float().cpu().numpy(), want[0]) <= (1e-4 if dtype == torch.float32 else 2e-2), fam


def test_selected_by_torch_deterministic_mode():
    inp = synthetic.make_inputs(2, 300, 4, 64, 4, [(1, 30), (1, 15)], torch.float32, DEV, seed=7, dist="uniform",
                                value="normal")
    ref = bwd(inp, "zeros", DET)[0]
    prev = torch.are_deterministic_algorithms_enabled()
    torch.use_deterministic_algorithms(True)
    try:
        grads = []
        for _ in range(2):
            v = inp["value"].detach().clone().requires_grad_(True)
            out = ms_deform_attn(v, inp["shapes"], inp["lsi"], inp["loc"], inp["aw"], 64)
            out.backward(inp["grad_out"].view_as(out))
            grads.append(v.grad)
    finally:
        torch.use_deterministic_algorithms(prev)
    assert torch.equal(grads[0], grads[1]) and torch.equal(grads[0], ref)


@pytest.mark.parametrize("scale", [1e-20, 1.0, 1e20])
def test_scale_follows_the_gradient_magnitude(oracle_c, scale):
    inp = synthetic.make_inputs(2, 41, 4, 32, 4, [(1, 20), (1, 10)], torch.float32, DEV, seed=8, dist="spill",
                                value="normal")
    inp["grad_out"] = inp["grad_out"] * scale
    f = {k: v for k, v in inp.items()}
    want = oracle_c.backward(*[f[k] for k in ("value", "shapes", "lsi", "loc", "aw")], f["grad_out"], "zeros")
    got = bwd(inp, "zeros", DET)
    assert rel_err(got[0].cpu().numpy(), want[0]) <= 1e-4
    zero = dict(inp)
    zero["grad_out"] = torch.zeros_like(inp["grad_out"])
    assert float(bwd(zero, "zeros", DET)[0].abs().max()) == 0.0


def test_deterministic_edge_cases():
    """(a) a non-finite grad_output under the deterministic flag surfaces as NaN in grad_value
    instead of finite garbage; (b) no value rows at all (S == 0) with the flag set does not touch
    the (absent) workspace."""
    inp = synthetic.make_inputs(2, 37, 4, 64, 4, [(1, 20), (1, 10)], torch.float32, DEV, seed=11, dist="uniform",
                                value="normal")
    inp["grad_out"][1, 3, 5] = float("inf")
    gv, gl, ga = bwd(inp, "zeros", DET)
    assert torch.isnan(gv).all()
    inp0 = synthetic.make_inputs(2, 5, 2, 64, 4, [(1, 0), (1, 0)], torch.float32, DEV, seed=12, dist="uniform")
    assert inp0["value"].shape[1] == 0
    gv, gl, ga = bwd(inp0, "zeros", DET)
    assert gv.numel() == 0 and float(gl.abs().max()) == 0.0 and float(ga.abs().max()) == 0.0
