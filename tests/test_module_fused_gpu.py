"""The fused module op (msda_module_forward / backward: softmax + sampling locations inside the
tile kernels) against the two-kernel path (msdeform_frontend +
ms_deform_attn), forward and all gradients; fp32 and bf16, both paddings, reference points with 1
and 2 columns, a ragged query count, offsets so wide that tiles fall back to the per-tap body, and
the drop-in module end to end (fused_op on / off)."""
import pytest
import torch

pytestmark = pytest.mark.gpu

import mfl_b200  # noqa: E402
from mfl_b200.functions import module_func, ms_deform_attn, msdeform_frontend  # noqa: E402
from mfl_b200 import synthetic  # noqa: E402

DEV = "cuda"


def rel(a, b):
    return float((a.double() - b.double()).abs().max() / b.double().abs().max().clamp_min(1e-30))


def make(N, T, M, D, dtype, ref_dim, seed, spread=1.0, Lq=None):
    g = torch.Generator(device=DEV).manual_seed(seed)
    L, P, S = len(T), 4, sum(T)
    Lq = S if Lq is None else Lq
    value = torch.randn(N, S, M, D, device=DEV, generator=g)
    ref0 = torch.rand(N, Lq, 1, 1, device=DEV, generator=g).expand(-1, -1, L, -1)
    if ref_dim == 2:
        ref = torch.cat((ref0, 0.05 + 0.1 * torch.rand(N, Lq, L, 1, device=DEV, generator=g)), -1)
    else:
        ref = ref0.contiguous()
    off = (synthetic.head_bias(M, P, DEV)[None, None, :, None, :] +
           spread * torch.randn(N, Lq, M, L, P, device=DEV, generator=g))
    logits = torch.randn(N, Lq, M, L * P, device=DEV, generator=g)
    gout = torch.randn(N, Lq, M * D, device=DEV, generator=g)
    Tt = torch.tensor(T, device=DEV)
    shapes = torch.stack((torch.ones_like(Tt), Tt), -1)
    lsi = torch.cat((Tt.new_zeros(1), Tt.cumsum(0)[:-1]))
    return dict(value=value.to(dtype), off=off.to(dtype), logits=logits.to(dtype), ref=ref.to(dtype).contiguous(),
                gout=gout.to(dtype), T=Tt, shapes=shapes, lsi=lsi)


def run(fused, c, padding):
    v, off, lg, ref = (c[k].detach().clone().requires_grad_(True) for k in ("value", "off", "logits", "ref"))
    if fused:
        out = module_func.msdeform_module(v, c["shapes"], c["lsi"], off, lg, ref, padding)
    else:
        loc, aw = msdeform_frontend(off, lg, ref, c["T"])
        out = ms_deform_attn(v, c["shapes"], c["lsi"], loc, aw, 64, padding)
    out.backward(c["gout"])
    return [t.float() for t in (out.detach(), v.grad, off.grad, lg.grad, ref.grad)]


CASES = [
    ("enc300", 2, [160, 80, 40, 20], 8, 64, 1, 1.0, None),
    ("ragged", 3, [100, 50, 25], 4, 64, 1, 0.5, 263),
    ("centre_length", 2, [128, 64, 32, 16], 8, 64, 2, 0.5, 300),
    ("d32", 2, [160, 80, 40, 20], 4, 32, 1, 0.5, None),
    ("wide_offsets_fall_back", 2, [160, 80, 40, 20], 8, 64, 1, 40.0, None),
]


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
@pytest.mark.parametrize("padding", ["zeros", "border"])
def test_fused_module_op_fp32(case, padding):
    _, N, T, M, D, rd, spread, Lq = case
    c = make(N, T, M, D, torch.float32, rd, seed=31, spread=spread, Lq=Lq)
    assert module_func.supported(c["value"], c["off"], c["ref"])
    got, want = run(True, c, padding), run(False, c, padding)
    assert float((got[0] - want[0]).abs().max()) <= 2e-5 * float(want[0].abs().max())
    for g_, w_, nm in zip(got[1:], want[1:], ("grad_value", "grad_offsets", "grad_logits", "grad_ref")):
        assert rel(g_, w_) <= 2e-4, nm


@pytest.mark.parametrize("padding", ["zeros", "border"])
def test_fused_module_op_bf16(padding):
    c = make(2, [160, 80, 40, 20], 8, 64, torch.bfloat16, 1, seed=32, spread=0.5)
    assert module_func.supported(c["value"], c["off"], c["ref"])
    got = run(True, c, padding)
    c32 = {k: (v.float() if v.is_floating_point() else v) for k, v in c.items()}
    want = run(False, c32, padding)                 # fp32 kernels on the bf16-rounded inputs
    assert rel(got[0], want[0]) <= 2.0 ** -7
    for g_, w_, nm in zip(got[1:], want[1:], ("grad_value", "grad_offsets", "grad_logits", "grad_ref")):
        assert rel(g_, w_) <= 2e-2, nm


@pytest.mark.parametrize("spread", [0.5, 40.0], ids=["tiles", "per_tap_fallback"])
def test_fused_module_op_padding_mask(spread):
    """The padding mask inside the kernels == value.masked_fill(mask, 0) in front of the two-kernel path,
    including the (zero) gradient of the masked value rows."""
    c = make(2, [160, 80, 40, 20], 8, 64, torch.float32, 1, seed=35, spread=spread)
    mask = torch.zeros(2, 300, dtype=torch.bool, device=DEV)
    mask[0, 120:160] = True
    mask[1, 20:60] = True
    mask[1, 290:] = True
    v, off, lg, ref = (c[k].detach().clone().requires_grad_(True) for k in ("value", "off", "logits", "ref"))
    out = module_func.msdeform_module(v, c["shapes"], c["lsi"], off, lg, ref, "zeros", mask)
    out.backward(c["gout"])
    v2, off2, lg2, ref2 = (c[k].detach().clone().requires_grad_(True) for k in ("value", "off", "logits", "ref"))
    loc, aw = msdeform_frontend(off2, lg2, ref2, c["T"])
    want = ms_deform_attn(v2.masked_fill(mask[..., None, None], 0.0), c["shapes"], c["lsi"], loc, aw, 64, "zeros")
    want.backward(c["gout"])
    assert float((out - want).abs().max()) <= 2e-5 * float(want.abs().max())
    assert float(v.grad[mask].abs().max()) == 0.0
    for g_, w_, nm in ((v.grad, v2.grad, "grad_value"), (off.grad, off2.grad, "grad_offsets"),
                       (lg.grad, lg2.grad, "grad_logits"), (ref.grad, ref2.grad, "grad_ref")):
        assert rel(g_, w_) <= 2e-4, nm


def test_not_supported_shapes_are_reported():
    c = make(2, [40, 20], 4, 64, torch.float32, 1, seed=33, Lq=37)          # decoder-like: short query list
    assert not module_func.supported(c["value"], c["off"], c["ref"])
    c = make(1, [160, 80], 4, 64, torch.float64, 1, seed=34)
    assert not module_func.supported(c["value"], c["off"], c["ref"])


def test_module_with_and_without_fused_op():
    torch.manual_seed(0)
    T = [160, 80, 40, 20]
    m = mfl_b200.MSDeformAttn(d_model=256, n_levels=4, n_heads=8, n_points=4).to(DEV)
    with torch.no_grad():
        m.attention_weights.weight.normal_(0, 0.05)
        m.sampling_offsets.weight.normal_(0, 0.01)
    N, S = 2, sum(T)
    shapes = torch.tensor(T, device=DEV)
    lsi = torch.cat((shapes.new_zeros(1), shapes.cumsum(0)[:-1]))
    x = torch.randn(N, S, 256, device=DEV)
    refp = synthetic.encoder_reference_points(T, DEV)[None, :, None, None].expand(N, -1, 4, 1).contiguous()
    mask = torch.zeros(N, S, dtype=torch.bool, device=DEV)
    mask[1, 250:] = True
    outs = []
    for fused in (True, False):
        m.fused_op = fused
        m.zero_grad()
        xi = x.clone().requires_grad_(True)
        out = m(xi, refp, xi, shapes, lsi, mask)
        out.square().sum().backward()
        outs.append((out.detach(), xi.grad.clone(), m.sampling_offsets.weight.grad.clone(),
                     m.attention_weights.weight.grad.clone()))
    for a, b in zip(*outs):
        torch.testing.assert_close(a, b, rtol=2e-3, atol=2e-4 * float(b.abs().max()))
    import pickle
    pickle.loads(pickle.dumps(m.cpu()))          # the caches do not travel


@pytest.mark.parametrize("spread", [0.5, 6.0], ids=["tiles", "some_fallback"])
def test_fused_module_op_saved_plan_on_and_off(spread):
    """The backward of the fused op from the forward's saved tile plans
    (MultiScaleDeformableAttention.SAVE_PLAN on) against the same op with it off: grad_offsets / grad_logits bitwise (same
    merged weights, same dots), grad_value / grad_ref up to the order of their atomics."""
    from mfl_b200 import MultiScaleDeformableAttention as MSDA
    if not MSDA.saved_plans_supported():
        pytest.skip("libmsda_b200.so built without -DMSDA_TILE_SAVED_PLANS=1 (the default: DESIGN.md section 6, 5h)")
    c = make(2, [160, 80, 40, 20], 8, 64, torch.float32, 2, seed=41, spread=spread, Lq=277)
    keep = MSDA.SAVE_PLAN
    try:
        MSDA.SAVE_PLAN = True
        a = run(True, c, "zeros")
        MSDA.SAVE_PLAN = False
        b = run(True, c, "zeros")
    finally:
        MSDA.SAVE_PLAN = keep
    assert torch.equal(a[0], b[0])
    assert torch.equal(a[2], b[2]) and torch.equal(a[3], b[3])
    assert rel(a[1], b[1]) <= 1e-5 and rel(a[4], b[4]) <= 1e-5
