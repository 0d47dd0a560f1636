"""GPU suite: the sm_100a kernels (called through the C ABI via the reference's
Python surface) against the oracles.

Tolerances (BASELINE.md section 5 / north_star):
  fp32 forward  max-abs <= 1e-5 on value in [0, 0.01)   (scaled by max|value| otherwise)
  fp32 grads    <= 1e-4 relative (to the tensor's max magnitude; atomics => never bitwise)
  fp64          1e-9-class; torch.autograd.gradcheck defaults
  bf16          vs the fp32 oracle on bf16-rounded inputs: forward rel <= 2^-7 (8e-3), grads rel <= 2e-2
"""
import numpy as np
import pytest
import torch

from conftest import GOLDEN_CASES, load_golden

pytestmark = pytest.mark.gpu

import mfl_b200  # noqa: E402
from mfl_b200 import MultiScaleDeformableAttention as MSDA  # noqa: E402
from mfl_b200 import _lib, synthetic  # noqa: E402
from mfl_b200.functions.ms_deform_attn_func import (  # noqa: E402
    MSDeformAttnBorderFunction, MSDeformAttnFunction, ms_deform_attn)

DEV = "cuda"
FWD_ABS_F32 = 1e-5
GRAD_REL_F32 = 1e-4
BF16_FWD_REL = 2.0 ** -7
BF16_GRAD_REL = 2e-2


def rel_err(a, b):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    denom = max(np.abs(b).max(), 1e-30) if b.size else 1.0
    return (np.abs(a - b).max() / denom) if b.size else 0.0


def run_cuda(inp, padding, flags=0):
    v, s, l, loc, aw, go = (inp[k] for k in ("value", "shapes", "lsi", "loc", "aw", "grad_out"))
    out = MSDA.ms_deform_attn_forward(v, s, l, loc, aw, 64, padding_mode=padding, flags=flags)
    gv, gl, ga = MSDA.ms_deform_attn_backward(v, s, l, loc, aw, go, 64, padding_mode=padding, flags=flags)
    torch.cuda.synchronize()
    return [t.float().cpu().numpy() if t.dtype == torch.bfloat16 else t.cpu().numpy()
            for t in (out, gv, gl, ga)]


def run_oracle(oracle_c, inp, padding):
    # bf16 inputs are rounded already; the oracle then works in fp32 on them
    args = [inp[k] for k in ("value", "shapes", "lsi", "loc", "aw")]
    out = oracle_c.forward(*args, padding)
    gv, gl, ga = oracle_c.backward(*args, inp["grad_out"], padding)
    return out, gv, gl, ga


# ---------------------------------------------------------------- golden fixtures
@pytest.mark.parametrize("name", GOLDEN_CASES)
@pytest.mark.parametrize("padding", ["zeros", "border"])
@pytest.mark.parametrize("flags", [0, _lib.MSDA_FLAG_FORCE_GENERIC])
def test_golden_from_reference(name, padding, flags):
    g = load_golden(name)
    inp = {k: torch.from_numpy(g[k]).to(DEV) for k in ("value", "shapes", "lsi", "loc", "aw", "grad_out")}
    out, gv, gl, ga = run_cuda(inp, padding, flags)
    if name.endswith("f64"):
        for got, key in ((out, "out_"), (gv, "gvalue_"), (gl, "gloc_"), (ga, "gaw_")):
            np.testing.assert_allclose(got, g[key + padding], rtol=1e-9, atol=1e-12)
    else:
        assert np.abs(out - g["out_" + padding]).max() <= FWD_ABS_F32
        for got, key in ((gv, "gvalue_"), (gl, "gloc_"), (ga, "gaw_")):
            assert rel_err(got, g[key + padding]) <= GRAD_REL_F32, key


# ---------------------------------------------------------------- sweep vs the C oracle
SWEEP = [
    # N, Lq, M, D, P, shapes
    (1, 2, 2, 2, 2, [(6, 4), (3, 2)]),                        # models/ops/test.py tiny
    (2, 37, 8, 64, 4, [(1, 38), (1, 19), (1, 10), (1, 5)]),    # production head size, temporal
    (2, 50, 8, 64, 4, [(7, 9), (4, 5), (2, 3), (1, 2)]),       # production head size, 2-D
    (3, 19, 4, 32, 1, [(1, 50)]),                              # L=1, P=1
    (2, 23, 2, 128, 8, [(1, 40), (1, 20), (1, 10), (1, 5)]),   # P=8 -> 32 points > lanes per group
    (2, 11, 3, 16, 3, [(5, 5), (3, 3), (2, 2)]),               # L*P = 9 (ragged point chunk)
    (1, 13, 2, 4, 2, [(1, 9), (1, 4)]),                        # one lane per head
    (2, 9, 2, 256, 2, [(3, 4), (2, 2)]),                       # D=256: generic f32, fast bf16
    (2, 7, 5, 30, 4, [(1, 20), (1, 10)]),                      # odd D -> generic
    (1, 5, 1, 71, 2, [(4, 4), (2, 2)]),                        # odd D -> generic
]


@pytest.mark.parametrize("case", SWEEP, ids=[f"N{c[0]}_Lq{c[1]}_M{c[2]}_D{c[3]}_P{c[4]}_L{len(c[5])}" for c in SWEEP])
@pytest.mark.parametrize("padding", ["zeros", "border"])
@pytest.mark.parametrize("dist", ["uniform", "spill"])
def test_fp32_vs_oracle(oracle_c, case, padding, dist):
    N, Lq, M, D, P, shapes = case
    inp = synthetic.make_inputs(N, Lq, M, D, P, shapes, torch.float32, DEV, seed=7, dist=dist)
    want = run_oracle(oracle_c, inp, padding)
    for flags in (0, _lib.MSDA_FLAG_FORCE_GENERIC, _lib.MSDA_FLAG_NO_WINDOW, _lib.MSDA_FLAG_FORCE_WINDOW):
        got = run_cuda(inp, padding, flags)
        assert np.abs(got[0] - want[0]).max() <= FWD_ABS_F32
        for g_, w_, nm in zip(got[1:], want[1:], ("grad_value", "grad_loc", "grad_aw")):
            assert rel_err(g_, w_) <= GRAD_REL_F32, (nm, flags)


@pytest.mark.parametrize("case", SWEEP[:6], ids=[f"N{c[0]}_Lq{c[1]}_M{c[2]}_D{c[3]}_P{c[4]}_L{len(c[5])}" for c in SWEEP[:6]])
@pytest.mark.parametrize("padding", ["zeros", "border"])
def test_fp64_vs_oracle(oracle_c, case, padding):
    N, Lq, M, D, P, shapes = case
    inp = synthetic.make_inputs(N, Lq, M, D, P, shapes, torch.float64, DEV, seed=9, dist="spill")
    got, want = run_cuda(inp, padding), run_oracle(oracle_c, inp, padding)
    for g_, w_ in zip(got, want):
        np.testing.assert_allclose(g_, w_, rtol=1e-9, atol=1e-13)


@pytest.mark.parametrize("case", SWEEP, ids=[f"N{c[0]}_Lq{c[1]}_M{c[2]}_D{c[3]}_P{c[4]}_L{len(c[5])}" for c in SWEEP])
@pytest.mark.parametrize("padding", ["zeros", "border"])
def test_bf16_vs_fp32_oracle_on_rounded_inputs(oracle_c, case, padding):
    N, Lq, M, D, P, shapes = case
    inp = synthetic.make_inputs(N, Lq, M, D, P, shapes, torch.bfloat16, DEV, seed=5, dist="uniform",
                                value="normal")
    want = run_oracle(oracle_c, {k: (v.float() if v.is_floating_point() else v) for k, v in inp.items()},
                      padding)
    for flags in (0, _lib.MSDA_FLAG_FORCE_GENERIC, _lib.MSDA_FLAG_FORCE_WINDOW):
        got = run_cuda(inp, padding, flags)
        assert rel_err(got[0], want[0]) <= BF16_FWD_REL
        for g_, w_, nm in zip(got[1:], want[1:], ("grad_value", "grad_loc", "grad_aw")):
            assert rel_err(g_, w_) <= BF16_GRAD_REL, (nm, flags)


# ---------------------------------------------------------------- the reference's own test script, as assertions
def _testpy_inputs(dtype, channels=2):
    # models/ops/test.py:17-35
    N, M, Lq, L, P = 1, 2, 2, 2, 2
    shapes = torch.as_tensor([(6, 4), (3, 2)], dtype=torch.long, device=DEV)
    lsi = torch.cat((shapes.new_zeros((1,)), shapes.prod(1).cumsum(0)[:-1]))
    S = int(shapes.prod(1).sum())
    torch.manual_seed(3)
    value = (torch.rand(N, S, M, channels, device=DEV) * 0.01).to(dtype)
    loc = torch.rand(N, Lq, M, L, P, 2, device=DEV).to(dtype)
    aw = torch.rand(N, Lq, M, L, P, device=DEV) + 1e-5
    aw = (aw / aw.sum(-1, keepdim=True).sum(-2, keepdim=True)).to(dtype)
    return value, shapes, lsi, loc, aw


@pytest.mark.parametrize("dtype,rtol,atol", [(torch.float64, 1e-5, 1e-8), (torch.float32, 1e-2, 1e-3)])
@pytest.mark.parametrize("padding", ["zeros", "border"])
def test_forward_equal_with_pytorch(dtype, rtol, atol, padding):
    """test.py:31-60 with its allclose tolerances; each padding mode against the
    grid_sample arrangement with the same mode."""
    from oracle.core_pytorch import ms_deform_attn_core_pytorch
    value, shapes, lsi, loc, aw = _testpy_inputs(dtype)
    want = ms_deform_attn_core_pytorch(value, shapes, loc, aw, padding).detach().cpu()
    got = ms_deform_attn(value, shapes, lsi, loc, aw, 2, padding).detach().cpu()
    assert torch.allclose(got, want, rtol=rtol, atol=atol)


@pytest.mark.parametrize("channels", [30, 32, 64, 71, 1025, 2048, 3096])
@pytest.mark.parametrize("fn", [MSDeformAttnFunction, MSDeformAttnBorderFunction])
def test_gradcheck_numerical(channels, fn):
    """test.py:63-86: gradcheck over one channel count per reference backward variant."""
    value, shapes, lsi, loc, aw = _testpy_inputs(torch.float64, channels)
    value.requires_grad_(True)
    loc.requires_grad_(True)
    aw.requires_grad_(True)
    assert torch.autograd.gradcheck(fn.apply, (value, shapes, lsi, loc, aw, 2))


# ---------------------------------------------------------------- boundary behaviour
def test_im2col_step_validation_and_independence():
    # cuda.cu:50-52: N % min(N, im2col_step) must be 0; results do not depend on the step
    shapes = [(1, 12), (1, 6)]
    inp = synthetic.make_inputs(96, 3, 2, 8, 2, shapes, torch.float32, DEV, seed=1)
    args = [inp[k] for k in ("value", "shapes", "lsi", "loc", "aw")]
    with pytest.raises(RuntimeError, match="im2col_step"):
        MSDeformAttnFunction.apply(*args, 64)
    a = MSDeformAttnFunction.apply(*args, 32)
    b = MSDeformAttnFunction.apply(*args, 96)
    assert torch.equal(a, b)
    inp = synthetic.make_inputs(128, 3, 2, 8, 2, shapes, torch.float32, DEV, seed=1)
    MSDeformAttnFunction.apply(*[inp[k] for k in ("value", "shapes", "lsi", "loc", "aw")], 64)


def test_contiguity_and_device_asserts():
    inp = synthetic.make_inputs(2, 3, 2, 8, 2, [(1, 12), (1, 6)], torch.float32, DEV, seed=1)
    v, s, l, loc, aw = (inp[k] for k in ("value", "shapes", "lsi", "loc", "aw"))
    with pytest.raises(RuntimeError, match="contiguous"):
        MSDeformAttnFunction.apply(v.transpose(2, 3), s, l, loc, aw, 64)
    with pytest.raises(RuntimeError, match="CUDA tensor"):
        MSDeformAttnFunction.apply(v, s.cpu(), l, loc, aw, 64)
    with pytest.raises(RuntimeError, match="Not implemented on the CPU"):
        MSDeformAttnFunction.apply(v.cpu(), s, l, loc, aw, 64)
    out = MSDeformAttnFunction.apply(v, s, l, loc, aw, 64)
    assert out.shape == (2, 3, 16)


def test_backward_return_pattern_and_nan_locations(oracle_c):
    inp = synthetic.make_inputs(2, 5, 2, 8, 2, [(1, 12), (1, 6)], torch.float32, DEV, seed=3,
                                requires_grad=True)
    with torch.no_grad():
        inp["loc"][0, 0, 0, 0, 0, 0] = float("nan")  # fails the gate: contributes 0, gets 0 grads
    out = MSDeformAttnFunction.apply(inp["value"], inp["shapes"], inp["lsi"], inp["loc"], inp["aw"], 64)
    out.backward(inp["grad_out"].view_as(out))
    assert torch.isfinite(out).all()
    assert inp["loc"].grad[0, 0, 0, 0, 0].abs().sum() == 0 and inp["aw"].grad[0, 0, 0, 0, 0] == 0
    assert inp["value"].grad.shape == inp["value"].shape


def test_empty_batch_and_queries():
    shapes, lsi, S = synthetic.level_tensors([(1, 4)], DEV)
    v = torch.rand(2, S, 2, 8, device=DEV)
    out = MSDA.ms_deform_attn_forward(v, shapes, lsi, torch.rand(2, 0, 2, 1, 2, 2, device=DEV),
                                      torch.rand(2, 0, 2, 1, 2, device=DEV), 64)
    assert out.shape == (2, 0, 16)
    gv, gl, ga = MSDA.ms_deform_attn_backward(v, shapes, lsi, torch.rand(2, 0, 2, 1, 2, 2, device=DEV),
                                              torch.rand(2, 0, 2, 1, 2, device=DEV),
                                              torch.rand(2, 0, 16, device=DEV), 64)
    assert float(gv.abs().sum()) == 0 and gl.numel() == 0 and ga.numel() == 0


def test_non_default_stream():
    inp = synthetic.make_inputs(2, 33, 8, 64, 4, [(1, 38), (1, 19), (1, 10), (1, 5)], torch.float32, DEV, seed=2)
    args = [inp[k] for k in ("value", "shapes", "lsi", "loc", "aw")]
    ref = MSDeformAttnFunction.apply(*args, 64)
    st = torch.cuda.Stream()
    st.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(st):
        out = MSDeformAttnFunction.apply(*args, 64)
    st.synchronize()
    assert torch.equal(out, ref)


def test_core_pytorch_name_runs_border_kernels():
    """functions.ms_deform_attn_core_pytorch (the name the reference exports,
    models/ops/functions/__init__.py:9) == the oracle port of the reference function, forward
    and gradients, on CUDA tensors."""
    from mfl_b200.functions import ms_deform_attn_core_pytorch
    from oracle.core_pytorch import ms_deform_attn_core_pytorch as ref_fn
    shapes_hw = [(6, 4), (3, 2)]
    inp = synthetic.make_inputs(2, 9, 2, 8, 2, shapes_hw, torch.float32, "cuda", seed=21, dist="spill",
                                requires_grad=True)
    out = ms_deform_attn_core_pytorch(inp["value"], shapes_hw, inp["loc"], inp["aw"])
    out.backward(inp["grad_out"])
    cpu = {k: v.detach().cpu() for k, v in inp.items()}
    v, loc, aw = (cpu[k].clone().requires_grad_(True) for k in ("value", "loc", "aw"))
    want = ref_fn(v, cpu["shapes"], loc, aw, "border")
    want.backward(cpu["grad_out"])
    assert float((out.detach().cpu() - want.detach()).abs().max()) <= 1e-5
    for got, ref in ((inp["value"].grad, v.grad), (inp["aw"].grad, aw.grad)):
        assert float((got.cpu() - ref).abs().max() / ref.abs().max().clamp_min(1e-30)) <= 1e-4
