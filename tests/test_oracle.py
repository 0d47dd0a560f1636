"""CPU suite: pin the oracles (oracle/msda_oracle.c and oracle/core_pytorch.py)
against the golden vectors produced by the reference's own Python
implementation (tests/golden/make_golden.py)."""
import numpy as np
import pytest
import torch

from conftest import GOLDEN_CASES, load_golden
from oracle import core_pytorch


def _tol(name):
    # fp64 goldens: the C restatement sums in a different order than
    # grid_sample -> tiny rounding differences only.  fp32: 1e-5-class.
    return dict(rtol=1e-9, atol=1e-12) if name.endswith("f64") else dict(rtol=2e-4, atol=2e-6)


@pytest.mark.parametrize("name", GOLDEN_CASES)
@pytest.mark.parametrize("padding", ["border", "zeros"])
def test_c_oracle_matches_reference_golden(oracle_c, name, padding):
    g = load_golden(name)
    out = oracle_c.forward(g["value"], g["shapes"], g["lsi"], g["loc"], g["aw"], padding)
    np.testing.assert_allclose(out, g["out_" + padding], **_tol(name))
    gv, gl, ga = oracle_c.backward(g["value"], g["shapes"], g["lsi"], g["loc"], g["aw"],
                                   g["grad_out"], padding)
    np.testing.assert_allclose(gv, g["gvalue_" + padding], **_tol(name))
    np.testing.assert_allclose(ga, g["gaw_" + padding], **_tol(name))
    np.testing.assert_allclose(gl[..., 0], g["gloc_" + padding][..., 0], **_tol(name))
    np.testing.assert_allclose(gl[..., 1], g["gloc_" + padding][..., 1], **_tol(name))


@pytest.mark.parametrize("name", ["temporal_f64", "temporal_f32"])
def test_live_twin_agrees(oracle_c, name):
    """the 1-D live twin (models/modules/attention.py:331-383) == 2-D form with y = 0.5"""
    g = load_golden(name)
    out = oracle_c.forward(g["value"], g["shapes"], g["lsi"], g["loc"], g["aw"], "border")
    np.testing.assert_allclose(out, g["out_live"], **_tol(name))
    gv, gl, ga = oracle_c.backward(g["value"], g["shapes"], g["lsi"], g["loc"], g["aw"],
                                   g["grad_out"], "border")
    np.testing.assert_allclose(gv, g["gvalue_live"], **_tol(name))
    np.testing.assert_allclose(ga, g["gaw_live"], **_tol(name))
    np.testing.assert_allclose(gl[..., 0], g["gloc_live"][..., 0], **_tol(name))


@pytest.mark.parametrize("name", GOLDEN_CASES)
@pytest.mark.parametrize("padding", ["border", "zeros"])
def test_torch_port_matches_reference_golden(name, padding):
    g = {k: torch.from_numpy(v) for k, v in load_golden(name).items()}
    value = g["value"].clone().requires_grad_(True)
    loc = g["loc"].clone().requires_grad_(True)
    aw = g["aw"].clone().requires_grad_(True)
    out = core_pytorch.ms_deform_attn_core_pytorch(value, g["shapes"], loc, aw, padding)
    out.backward(g["grad_out"])
    tol = _tol(name)
    torch.testing.assert_close(out.detach(), g["out_" + padding], **tol)
    torch.testing.assert_close(value.grad, g["gvalue_" + padding], **tol)
    torch.testing.assert_close(loc.grad, g["gloc_" + padding], **tol)
    torch.testing.assert_close(aw.grad, g["gaw_" + padding], **tol)


def test_interior_zeros_equals_border():
    g = load_golden("interior2d_f32")
    np.testing.assert_array_equal(g["out_zeros"], g["out_border"])
