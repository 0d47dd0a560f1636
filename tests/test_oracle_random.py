"""CPU suite: the two oracles against each other -- and, in the build container, against the
reference's own function imported unmodified from /root/reference -- on seeded random calls far
from the six golden fixtures: ragged level lists (1-row and 2-D maps, single-token levels),
L and P of 1, odd head sizes, sampling locations well outside [0, 1], no queries at all.

oracle/msda_oracle.c restates the CUDA kernels (models/ops/src/cuda/ms_deform_im2col_cuda.cuh
:34-160, 238-300), oracle/core_pytorch.py the grid_sample path (models/ops/functions/
ms_deform_attn_func.py:44-71): two independent restatements of the same operator, so agreement
to fp64 rounding on inputs neither was written against is what "the oracle is pinned" rests on
beyond the fixtures.  /root/reference does not exist on the GPU box: that leg skips there.
"""
import importlib.util
import os

import numpy as np
import pytest
import torch

from oracle import core_pytorch

REF_FUNC = "/root/reference/models/ops/functions/ms_deform_attn_func.py"

# (N, M, D, levels as (H, W), Lq, P)
CASES = [
    (2, 8, 32, [(1, 37), (1, 19), (1, 10), (1, 5)], 23, 4),      # the repo's temporal layout
    (1, 3, 7, [(1, 1), (1, 2), (1, 64)], 9, 3),                  # single-token level, odd D
    (2, 2, 5, [(6, 4), (3, 2)], 11, 2),                          # 2-D maps (test.py:17-44 class)
    (3, 1, 16, [(1, 50)], 17, 1),                                # L = 1, P = 1, one head
    (1, 4, 4, [(2, 3), (1, 7), (5, 1)], 5, 5),                   # mixed, a 1-column map
    (2, 2, 8, [(1, 12), (1, 6)], 0, 4),                          # no queries
]


def make_case(case, seed, spread):
    N, M, D, levels, Lq, P = case
    g = torch.Generator().manual_seed(seed)
    shapes = torch.tensor(levels, dtype=torch.long)
    sizes = shapes.prod(1)
    lsi = torch.cat((sizes.new_zeros(1), sizes.cumsum(0)[:-1]))
    S, L = int(sizes.sum()), len(levels)
    value = torch.randn(N, S, M, D, dtype=torch.float64, generator=g)
    # spread 0: inside the maps; 0.6: a good share of the taps hang over an edge or miss the map
    loc = torch.rand(N, Lq, M, L, P, 2, dtype=torch.float64, generator=g) * (1 + 2 * spread) - spread
    aw = torch.rand(N, Lq, M, L, P, dtype=torch.float64, generator=g) + 1e-3
    aw = aw / aw.sum((-1, -2), keepdim=True).clamp_min(1e-30)
    grad_out = torch.randn(N, Lq, M * D, dtype=torch.float64, generator=g)
    return value, shapes, lsi, loc, aw, grad_out


def torch_port(value, shapes, loc, aw, grad_out, padding):
    v, l_, a = (t.clone().requires_grad_(True) for t in (value, loc, aw))
    out = core_pytorch.ms_deform_attn_core_pytorch(v, shapes, l_, a, padding)
    out.backward(grad_out)
    return out.detach(), v.grad, l_.grad, a.grad


def away_from_pixel_crossings(loc, shapes, eps=1e-6):
    """grad_sampling_loc jumps where a pixel coordinate is an integer; random fp64 inputs do not
    land there, but make the test say so instead of relying on it."""
    wh = torch.stack((shapes[:, 1], shapes[:, 0]), -1).to(loc.dtype)[None, None, None, :, None, :]
    pix = loc * wh - 0.5
    return bool(((pix - pix.round()).abs() > eps).all())


@pytest.mark.parametrize("padding", ["zeros", "border"])
@pytest.mark.parametrize("spread", [0.0, 0.6])
@pytest.mark.parametrize("case", range(len(CASES)))
def test_c_restatement_equals_grid_sample_port(oracle_c, case, spread, padding):
    value, shapes, lsi, loc, aw, grad_out = make_case(CASES[case], 100 + case, spread)
    assert away_from_pixel_crossings(loc, shapes)
    out_c = oracle_c.forward(value, shapes, lsi, loc, aw, padding)
    gv_c, gl_c, ga_c = oracle_c.backward(value, shapes, lsi, loc, aw, grad_out, padding)
    if loc.shape[1] == 0:
        # no query: an empty output, zero value gradient (ms_deform_attn_cuda.cu:121-123)
        assert out_c.shape == (value.shape[0], 0, value.shape[2] * value.shape[3])
        assert not gv_c.any() and gl_c.size == 0 and ga_c.size == 0
        return
    out_t, gv_t, gl_t, ga_t = torch_port(value, shapes, loc, aw, grad_out, padding)
    tol = dict(rtol=1e-9, atol=1e-11)
    np.testing.assert_allclose(out_c, out_t.numpy(), **tol)
    np.testing.assert_allclose(gv_c, gv_t.numpy(), **tol)
    np.testing.assert_allclose(ga_c, ga_t.numpy(), **tol)
    np.testing.assert_allclose(gl_c, gl_t.numpy(), **tol)


@pytest.mark.skipif(not os.path.exists(REF_FUNC), reason="/root/reference is only in the build container")
@pytest.mark.parametrize("spread", [0.0, 0.6])
@pytest.mark.parametrize("case", range(len(CASES) - 1))
def test_both_oracles_equal_the_live_reference_function(oracle_c, case, spread):
    """The reference's ms_deform_attn_core_pytorch, imported by path and run as it is (it
    hard-codes grid_sample's padding_mode='border', ms_deform_attn_func.py:61)."""
    spec = importlib.util.spec_from_file_location("_ref_msda_func_random", REF_FUNC)
    ref = importlib.util.module_from_spec(spec)
    try:
        spec.loader.exec_module(ref)        # imports the compiled extension at module level
    except ImportError:
        # same dodge as tests/golden/make_golden.py: the function itself needs only torch
        import sys
        import types
        sys.modules.setdefault("MultiScaleDeformableAttention", types.ModuleType("MultiScaleDeformableAttention"))
        spec.loader.exec_module(ref)
    value, shapes, lsi, loc, aw, grad_out = make_case(CASES[case], 200 + case, spread)
    v, l_, a = (t.clone().requires_grad_(True) for t in (value, loc, aw))
    out_r = ref.ms_deform_attn_core_pytorch(v, shapes, l_, a)
    out_r.backward(grad_out)
    out_t, gv_t, gl_t, ga_t = torch_port(value, shapes, loc, aw, grad_out, "border")
    # the port is the same sequence of torch calls: bit-for-bit
    assert torch.equal(out_r.detach(), out_t) and torch.equal(v.grad, gv_t)
    assert torch.equal(l_.grad, gl_t) and torch.equal(a.grad, ga_t)
    tol = dict(rtol=1e-9, atol=1e-11)
    np.testing.assert_allclose(oracle_c.forward(value, shapes, lsi, loc, aw, "border"), out_r.detach().numpy(), **tol)
    gv_c, gl_c, ga_c = oracle_c.backward(value, shapes, lsi, loc, aw, grad_out, "border")
    np.testing.assert_allclose(gv_c, v.grad.numpy(), **tol)
    np.testing.assert_allclose(gl_c, l_.grad.numpy(), **tol)
    np.testing.assert_allclose(ga_c, a.grad.numpy(), **tol)
