"""GPU parity of the fused module front-end (msda_frontend_forward / _backward, SURVEY.md
section 8(f) row 1) against the restatement of the module's own torch chain
(oracle/core_pytorch.py::frontend_pytorch, models/ops/modules/ms_deform_attn.py:98-117).

fp64: 1e-12; fp32: 1e-6 relative (expf vs torch's exp, different summation order);
bf16: outputs compared with the fp32 chain on the bf16-rounded inputs, 2^-7 relative.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu

import mfl_b200  # noqa: E402,F401
from mfl_b200.functions import msdeform_frontend  # noqa: E402
from mfl_b200.modules import MSDeformAttn  # noqa: E402
from oracle.core_pytorch import frontend_pytorch  # noqa: E402

DEV = "cuda"


def rel(a, b):
    return float((a.double() - b.double()).abs().max() / b.double().abs().max().clamp_min(1e-30))


def make(N, Lq, M, L, P, rd, dtype, seed=0, ref_grad=True):
    g = torch.Generator(device=DEV).manual_seed(seed)
    off = torch.randn(N, Lq, M, L, P, device=DEV, generator=g, dtype=torch.float64) * 2
    lg = torch.randn(N, Lq, M, L * P, device=DEV, generator=g, dtype=torch.float64) * 3
    ref = torch.rand(N, Lq, L, rd, device=DEV, generator=g, dtype=torch.float64)
    sizes = torch.tensor([max(3, 50 >> l) for l in range(L)], dtype=torch.long, device=DEV)
    gl = torch.randn(N, Lq, M, L, P, 2, device=DEV, generator=g, dtype=torch.float64)
    ga = torch.randn(N, Lq, M, L, P, device=DEV, generator=g, dtype=torch.float64)
    t = [x.to(dtype) for x in (off, lg, ref, gl, ga)]
    t[0].requires_grad_(True)
    t[1].requires_grad_(True)
    t[2].requires_grad_(ref_grad)
    return t[0], t[1], t[2], sizes, t[3], t[4]


def run(fn, off, lg, ref, sizes, gl, ga):
    for t in (off, lg, ref):
        t.grad = None
    loc, aw = fn(off, lg, ref, sizes)
    torch.autograd.backward((loc, aw), (gl, ga))
    return loc.detach(), aw.detach(), off.grad, lg.grad, ref.grad


CASES = [
    # N, Lq, M, L, P, ref_dim
    (2, 37, 8, 4, 4, 1),     # the production layout: 16-byte pieces, shuffle reduction over 8 heads
    (2, 37, 8, 4, 4, 2),     # (centre, length) reference boxes
    (1, 50, 3, 2, 4, 1),     # M = 3: grad_ref through atomics
    (2, 9, 4, 3, 3, 2),      # P = 3: element-wise path
    (1, 5, 1, 1, 8, 1),      # one head, one level, 8 points
    (3, 130, 16, 4, 4, 1),   # 16 heads; 130*16 groups: a ragged last block
    (1, 7, 32, 2, 2, 2),     # a full warp of heads
]
IDS = ["prod", "prod_box", "M3_atomic", "P3", "M1L1P8", "M16", "M32"]


@pytest.mark.parametrize("case", CASES, ids=IDS)
@pytest.mark.parametrize("dtype,tol", [(torch.float64, 1e-12), (torch.float32, 2e-6)], ids=["f64", "f32"])
def test_frontend_matches_torch_chain(case, dtype, tol):
    args = make(*case, dtype, seed=3)
    got = run(msdeform_frontend, *args)
    want = run(frontend_pytorch, *args)
    for g, w, name in zip(got, want, ("loc", "aw", "grad_offsets", "grad_logits", "grad_ref")):
        assert g.shape == w.shape, name
        assert rel(g, w) <= tol, name
    assert bool((got[0][..., 1] == 0.5).all())


@pytest.mark.parametrize("case", CASES[:4], ids=IDS[:4])
def test_frontend_bf16(case):
    args = make(*case, torch.bfloat16, seed=4)
    got = run(msdeform_frontend, *args)
    f32 = [a.detach().float().requires_grad_(a.requires_grad) if a.is_floating_point() else a for a in args]
    want = run(frontend_pytorch, *f32)
    for g, w, name in zip(got, want, ("loc", "aw", "grad_offsets", "grad_logits", "grad_ref")):
        assert g.dtype == torch.bfloat16, name
        # M3_atomic accumulates grad_ref in bf16 with atomics: 12 roundings instead of one
        assert rel(g.float(), w) <= (2.0 ** -7 if not (name == "grad_ref" and case[2] == 3) else 3e-2), name


def test_frontend_without_reference_gradient_and_empty():
    args = make(2, 11, 8, 4, 4, 1, torch.float32, seed=5, ref_grad=False)
    got = run(msdeform_frontend, *args)
    want = run(frontend_pytorch, *args)
    assert got[4] is None and want[4] is None
    assert rel(got[2], want[2]) <= 2e-6 and rel(got[3], want[3]) <= 2e-6
    e = make(0, 5, 8, 4, 4, 1, torch.float32)
    loc, aw = msdeform_frontend(e[0], e[1], e[2], e[3])
    assert loc.shape == (0, 5, 8, 4, 4, 2) and aw.shape == (0, 5, 8, 4, 4)


def test_frontend_gradcheck():
    off, lg, ref, sizes, _, _ = make(1, 3, 2, 2, 4, 2, torch.float64, seed=6)
    assert torch.autograd.gradcheck(lambda a, b, c: msdeform_frontend(a, b, c, sizes), (off, lg, ref),
                                    eps=1e-6, atol=1e-7, rtol=1e-5, nondet_tol=0.0)


@pytest.mark.parametrize("rd", [1, 2])
def test_module_fused_frontend_equals_torch_chain(rd):
    torch.manual_seed(0)
    mod = MSDeformAttn(256, 4, 8, 4, attn_weight_init="xavier").to(DEV)
    with torch.no_grad():
        mod.sampling_offsets.weight.normal_(0, 0.02)
    T = torch.tensor([48, 24, 12, 6], dtype=torch.long, device=DEV)
    lsi = torch.cat((T.new_zeros(1), T.cumsum(0)[:-1]))
    N, Lq, S = 2, 33, int(T.sum())
    q = torch.randn(N, Lq, 256, device=DEV, requires_grad=True)
    src = torch.randn(N, S, 256, device=DEV, requires_grad=True)
    ref = torch.rand(N, Lq, 4, rd, device=DEV, requires_grad=True)
    res = []
    for fused in (True, False):
        mod.fused_frontend = fused
        for t in (q, src, ref):
            t.grad = None
        mod.zero_grad()
        out, loc, aw = mod(q, ref, src, T, lsi, is_sparse=True)
        out.square().mean().backward()
        res.append([out.detach(), loc.detach(), aw.detach(), q.grad.clone(), src.grad.clone(), ref.grad.clone(),
                    mod.sampling_offsets.weight.grad.clone(), mod.attention_weights.weight.grad.clone()])
    assert res[0][1].shape == (N, Lq, 8, 4, 4, 1)
    for a, b in zip(*res):
        assert rel(a, b) <= 1e-4
