"""GPU suite for the row-window kernels (csrc/msda_win.cuh) and the CTA-tile kernels
(csrc/msda_tile.cuh: time-sorted query tiles, TMA-filled shared memory) against the C oracle.
Every case runs through the families (`fam`): "tile" is what the shape class gets by default,
"window" is the same call with MSDA_FLAG_NO_TILE, "plan" the forward as plan kernel + persistent
apply kernel (MSDA_FLAG_PLAN, csrc/msda_plan.cuh; the backward of that family is the tile kernel).

The window path is taken per block of 4 consecutive queries whose packed per-level
windows fit kWinCap rows; everything else in the same launch runs the per-tap body.
These cases force the window kernels (MSDA_FLAG_FORCE_WINDOW) at sizes the oracle
finishes in seconds and cover: locality (model-like locations), ragged query counts
(Lq % 16 != 0, Lq < 4), 1..4 levels, both paddings, vertical coordinates != 0.5 on
(1, W) maps, taps that fall off either end of a level, NaN locations, blocks that
straddle a level boundary (fallback inside a window launch) and bf16.

Tolerances as in tests/test_parity_gpu.py.
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

import mfl_b200  # noqa: E402,F401
from mfl_b200 import MultiScaleDeformableAttention as MSDA  # noqa: E402
from mfl_b200 import _lib, synthetic  # noqa: E402

DEV = "cuda"
FWD_ABS_F32 = 1e-5
GRAD_REL_F32 = 1e-4
BF16_FWD_REL = 2.0 ** -7
BF16_GRAD_REL = 2e-2
WIN = _lib.MSDA_FLAG_FORCE_WINDOW
SAVE = _lib.MSDA_FLAG_SAVE_PLAN
FAMILY_FLAGS = {"tile": WIN, "window": WIN | _lib.MSDA_FLAG_NO_TILE, "plan": WIN | _lib.MSDA_FLAG_PLAN,
                "saved": WIN | SAVE,   # saved: the backward starts from the tile plans the forward saved
                "persist": WIN | _lib.MSDA_FLAG_PERSIST}   # the forward tile kernel in its persistent form


@pytest.fixture(params=["tile", "window", "plan", "saved", "persist"])
def fam(request):
    if request.param == "saved" and not MSDA.saved_plans_supported():
        pytest.skip("libmsda_b200.so built without -DMSDA_TILE_SAVED_PLANS=1 (the default: DESIGN.md section 6, 5h)")
    return FAMILY_FLAGS[request.param]


def rel_err(a, b):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return (np.abs(a - b).max() / max(np.abs(b).max(), 1e-30)) if b.size else 0.0


def run_cuda(inp, padding, flags):
    v, s, l, loc, aw, go = (inp[k] for k in ("value", "shapes", "lsi", "loc", "aw", "grad_out"))
    if flags & SAVE:
        out, plan = MSDA.ms_deform_attn_forward(v, s, l, loc, aw, 64, padding_mode=padding, flags=flags & ~SAVE,
                                                return_plan=True, save_plan=True)
        assert plan is None or plan.msda_saved
        gv, gl, ga = MSDA.ms_deform_attn_backward(v, s, l, loc, aw, go, 64, padding_mode=padding,
                                                  flags=flags & ~SAVE, plan=plan)
    else:
        out = MSDA.ms_deform_attn_forward(v, s, l, loc, aw, 64, padding_mode=padding, flags=flags)
        gv, gl, ga = MSDA.ms_deform_attn_backward(v, s, l, loc, aw, go, 64, padding_mode=padding, flags=flags)
    torch.cuda.synchronize()
    return [t.float().cpu().numpy() for t in (out, gv, gl, ga)]


def run_oracle(oracle_c, inp, padding):
    f = {k: (v.float() if v.is_floating_point() else v) for k, v in inp.items()}
    args = [f[k] for k in ("value", "shapes", "lsi", "loc", "aw")]
    out = oracle_c.forward(*args, padding)
    gv, gl, ga = oracle_c.backward(*args, f["grad_out"], padding)
    return out, gv, gl, ga


def check_f32(got, want, smooth=None):
    assert np.abs(got[0] - want[0]).max() <= FWD_ABS_F32
    assert rel_err(got[1], want[1]) <= GRAD_REL_F32, "grad_value"
    assert rel_err(got[3], want[3]) <= GRAD_REL_F32, "grad_attn_weight"
    gl_g, gl_w = got[2], want[2]
    if smooth is not None:
        gl_g, gl_w = gl_g * smooth, gl_w * smooth
    assert rel_err(gl_g, gl_w) <= GRAD_REL_F32, "grad_sampling_loc"


def smooth_mask(inp, eps=1e-4):
    """grad_sampling_loc jumps where a pixel coordinate crosses an integer; fp32 rounding of
    loc*W-0.5 may differ by an ulp between two correct implementations (tests/test_fullsize_gpu.py)."""
    loc = inp["loc"].double().cpu()
    sh = inp["shapes"].cpu()
    wh = torch.stack((sh[:, 1], sh[:, 0]), -1).double()[None, None, None, :, None, :]
    pix = loc * wh - 0.5
    return (((pix - pix.round()).abs() > eps) | (wh == 1)).all(-1, keepdim=True).numpy()


CASES = [
    # N, T (levels), M, D, enc (queries = tokens) or Lq
    (2, [96, 48, 24, 12], 8, 64, None),      # encoder-like, Lq = 180 (not a multiple of 16)
    (3, [50, 25], 4, 64, None),              # two levels, Lq = 75
    (2, [40], 2, 64, None),                  # one level
    (2, [33, 17, 9], 3, 64, None),           # three levels, odd sizes
    (2, [64, 32, 16, 8], 8, 32, None),       # D = 32: one 16-byte unit per lane
    (1, [20, 10, 5, 3], 2, 64, 3),           # fewer queries than one block
    (2, [80, 40, 20, 10], 4, 64, 37),        # decoder-like random reference points (mostly fallback)
]
IDS = ["enc180", "L2", "L1", "L3odd", "D32", "Lq3", "dec37"]


def make(case, dtype, seed, dist="model"):
    N, T, M, D, Lq = case
    enc = Lq is None
    return synthetic.make_inputs(N, sum(T) if enc else Lq, M, D, 4, [(1, t) for t in T], dtype, DEV, seed=seed,
                                 dist=dist, value="small" if dtype == torch.float32 else "normal", enc_ref=enc)


@pytest.mark.parametrize("case", CASES, ids=IDS)
@pytest.mark.parametrize("padding", ["zeros", "border"])
def test_window_fp32_model_locations(oracle_c, case, padding, fam):
    inp = make(case, torch.float32, seed=11)
    check_f32(run_cuda(inp, padding, fam), run_oracle(oracle_c, inp, padding), smooth_mask(inp))


@pytest.mark.parametrize("case", CASES[:4], ids=IDS[:4])
@pytest.mark.parametrize("padding", ["zeros", "border"])
def test_window_matches_per_tap_kernels(case, padding, fam):
    """Same launch arguments through both kernel families."""
    inp = make(case, torch.float32, seed=12)
    a = run_cuda(inp, padding, fam)
    b = run_cuda(inp, padding, _lib.MSDA_FLAG_NO_WINDOW)
    assert np.abs(a[0] - b[0]).max() <= 2e-6
    m = smooth_mask(inp)
    for x, y, nm in ((a[1], b[1], "gv"), (a[2] * m, b[2] * m, "gl"), (a[3], b[3], "ga")):
        assert rel_err(x, y) <= 2e-5, nm


@pytest.mark.parametrize("padding", ["zeros", "border"])
def test_window_vertical_coordinate_and_edges(oracle_c, padding, fam):
    """(1, W) maps with y != 0.5 (the single row is the low or the high row of the vertical
    pair), x running off both ends, and a NaN location."""
    inp = make(CASES[0], torch.float32, seed=13)
    g = torch.Generator(device=DEV).manual_seed(5)
    loc = inp["loc"].clone()
    loc[..., 1] = torch.rand(loc.shape[:-1], device=DEV, generator=g) * 2.4 - 0.7      # h_im in (-1.2, 1.2)
    loc[:, :8, :, :, :, 0] -= 0.02                                                      # off the left end
    loc[:, -8:, :, :, :, 0] += 0.05                                                     # off the right end
    loc[0, 17, 1, 2, 3, 0] = float("nan")
    inp["loc"] = loc.contiguous()
    got = run_cuda(inp, padding, fam)
    assert np.isfinite(got[0]).all()
    ref = dict(inp)
    if padding == "zeros":
        # the oracle treats NaN like the reference: the gate fails, the point contributes nothing
        want = run_oracle(oracle_c, ref, padding)
        check_f32(got, want, smooth_mask(inp))
        assert np.abs(got[2][0, 17, 1, 2, 3]).sum() == 0 and got[3][0, 17, 1, 2, 3] == 0
    else:
        ref["loc"] = torch.nan_to_num(inp["loc"], nan=-1.0)   # border: a NaN x clamps to column 0
        want = run_oracle(oracle_c, ref, padding)
        m = smooth_mask(ref)
        m[0, 17, 1, 2, 3] = 0
        check_f32(got, want, m)


def test_window_uniform_locations_fall_back(oracle_c, fam):
    """Random locations: every block / tile exceeds the window capacity and runs the per-tap body
    inside the window / tile launch."""
    inp = make(CASES[0], torch.float32, seed=14, dist="uniform")
    check_f32(run_cuda(inp, "zeros", fam), run_oracle(oracle_c, inp, "zeros"), smooth_mask(inp))


@pytest.mark.parametrize("case", [CASES[0], CASES[1]], ids=[IDS[0], IDS[1]])
@pytest.mark.parametrize("padding", ["zeros", "border"])
def test_window_bf16(oracle_c, case, padding, fam):
    inp = make(case, torch.bfloat16, seed=15)
    got, want = run_cuda(inp, padding, fam), run_oracle(oracle_c, inp, padding)
    assert rel_err(got[0], want[0]) <= BF16_FWD_REL
    for g_, w_, nm in zip(got[1:], want[1:], ("grad_value", "grad_loc", "grad_aw")):
        assert rel_err(g_, w_) <= BF16_GRAD_REL, nm


def test_permuted_queries_tile_bitwise_equal():
    """The tile kernels take the queries in time order: permuting the query list (the sparse
    encoders' score-ordered top-k tokens) permutes the output rows and nothing else, bitwise."""
    inp = make(CASES[0], torch.float32, seed=17)
    a = [inp[k] for k in ("value", "shapes", "lsi", "loc", "aw")]
    o0 = MSDA.ms_deform_attn_forward(*a, 64, flags=WIN)
    g = torch.Generator(device=DEV).manual_seed(3)
    perm = torch.randperm(inp["loc"].shape[1], device=DEV, generator=g)
    a[3], a[4] = inp["loc"][:, perm].contiguous(), inp["aw"][:, perm].contiguous()
    o1 = MSDA.ms_deform_attn_forward(*a, 64, flags=WIN)
    assert torch.equal(o0[:, perm], o1)


def test_tile_2d_maps_and_forward_plan_reuse(oracle_c):
    """(a) maps with H != 1 inside a tile launch run the 4-tap body; (b) the plan msda_forward_ex
    leaves behind gives the same backward as a backward that sorts for itself."""
    inp = synthetic.make_inputs(2, 300, 4, 64, 4, [(4, 8), (2, 4)], torch.float32, DEV, seed=18, dist="uniform")
    check_f32(run_cuda(inp, "zeros", WIN), run_oracle(oracle_c, inp, "zeros"), smooth_mask(inp))
    inp = make(CASES[0], torch.float32, seed=19)
    v, s, l, loc, aw, go = (inp[k] for k in ("value", "shapes", "lsi", "loc", "aw", "grad_out"))
    out, plan = MSDA.ms_deform_attn_forward(v, s, l, loc, aw, 64, flags=WIN, return_plan=True)
    assert plan is not None
    srt = plan[:2 * loc.shape[1]].view(2, -1).sort(1)[0]
    assert torch.equal(srt, torch.arange(loc.shape[1], device=DEV, dtype=torch.int32).expand(2, -1))
    g0 = MSDA.ms_deform_attn_backward(v, s, l, loc, aw, go, 64, flags=WIN, plan=plan)
    g1 = MSDA.ms_deform_attn_backward(v, s, l, loc, aw, go, 64, flags=WIN)
    assert torch.equal(g0[1], g1[1]) and torch.equal(g0[2], g1[2])
    assert rel_err(g0[0].cpu().numpy(), g1[0].cpu().numpy()) <= 1e-6


def test_window_default_heuristic_and_determinism():
    """Lq >= 256 takes the window kernels by default; the forward is bitwise reproducible."""
    inp = synthetic.make_inputs(2, 480, 8, 64, 4, [(1, 256), (1, 128), (1, 64), (1, 32)], torch.float32, DEV,
                                seed=16, dist="model", enc_ref=True)
    a = [inp[k] for k in ("value", "shapes", "lsi", "loc", "aw")]
    o0 = MSDA.ms_deform_attn_forward(*a, 64)
    o1 = MSDA.ms_deform_attn_forward(*a, 64, flags=WIN)
    o2 = MSDA.ms_deform_attn_forward(*a, 64, flags=_lib.MSDA_FLAG_NO_WINDOW)
    assert torch.equal(o0, o1)
    assert float((o0 - o2).abs().max()) <= 2e-6 and not torch.equal(o0, o2)


def test_long_query_list_uses_single_cta_sort():
    """Lq > 4096: the query order comes from msda_sort_kernel (one CTA per batch item, histogram with
    atomics) instead of the cluster sort; same results as the per-tap kernels."""
    T = [2048, 1024, 512, 256]
    inp = synthetic.make_inputs(1, 5000, 2, 64, 4, [(1, t) for t in T], torch.float32, DEV, seed=23, dist="model",
                                value="small", enc_ref=False)
    # reference points monotone in the (shuffled) query index would be too easy: make_inputs draws them at random
    a = run_cuda(inp, "zeros", WIN)
    b = run_cuda(inp, "zeros", _lib.MSDA_FLAG_NO_WINDOW)
    assert np.abs(a[0] - b[0]).max() <= 2e-6
    m = smooth_mask(inp)
    for x, y, nm in ((a[1], b[1], "gv"), (a[2] * m, b[2] * m, "gl"), (a[3], b[3], "ga")):
        assert rel_err(x, y) <= 2e-5, nm
    out, plan = MSDA.ms_deform_attn_forward(*[inp[k] for k in ("value", "shapes", "lsi", "loc", "aw")], 64, flags=WIN,
                                            return_plan=True)
    srt = plan[:5000].sort()[0]
    assert torch.equal(srt, torch.arange(5000, device=DEV, dtype=torch.int32))


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("sigma", [0.5, 3.0])
def test_saved_plan_backward_equals_recomputed_plan(dtype, sigma):
    """MSDA_FLAG_SAVE_PLAN: the backward that starts from the forward's saved tile plans gives the
    gradients of the backward that rebuilds them from the points -- grad_sampling_loc and
    grad_attn_weight bitwise (same weights, same dots), grad_value up to the order of its atomics.
    sigma = 3 rows: part of the tiles have no plan (per-tap body) and say so in their header.
    Ragged query count (last tile 37 queries), plan buffer pre-filled with garbage."""
    if not MSDA.saved_plans_supported():
        pytest.skip("libmsda_b200.so built without -DMSDA_TILE_SAVED_PLANS=1 (the default: DESIGN.md section 6, 5h)")
    T = [256, 128, 64, 32]
    inp = synthetic.make_inputs(2, 421, 8, 64, 4, [(1, t) for t in T], dtype, DEV, seed=31, dist="model",
                                value="normal", sigma=sigma)
    v, s, l, loc, aw, go = (inp[k] for k in ("value", "shapes", "lsi", "loc", "aw", "grad_out"))
    junk = torch.full((1 << 22,), -1, dtype=torch.int32, device=DEV)   # whatever the allocator hands out next
    del junk
    o_s, plan_s = MSDA.ms_deform_attn_forward(v, s, l, loc, aw, 64, return_plan=True, save_plan=True)
    o_n, plan_n = MSDA.ms_deform_attn_forward(v, s, l, loc, aw, 64, return_plan=True, save_plan=False)
    assert plan_s.msda_saved and not plan_n.msda_saved and plan_s.numel() > plan_n.numel()
    assert torch.equal(o_s, o_n)
    nq = loc.shape[0] * loc.shape[1]
    assert torch.equal(plan_s[:nq], plan_n[:nq])
    g_s = MSDA.ms_deform_attn_backward(v, s, l, loc, aw, go, 64, plan=plan_s)
    g_n = MSDA.ms_deform_attn_backward(v, s, l, loc, aw, go, 64, plan=plan_n)
    assert torch.equal(g_s[1], g_n[1]) and torch.equal(g_s[2], g_n[2])
    tol = 1e-5 if dtype == torch.float32 else BF16_GRAD_REL
    assert rel_err(g_s[0].float().cpu().numpy(), g_n[0].float().cpu().numpy()) <= tol
    # and once more from the same plan (the backward leaves it untouched)
    g_r = MSDA.ms_deform_attn_backward(v, s, l, loc, aw, go, 64, plan=plan_s)
    assert torch.equal(g_r[1], g_s[1]) and torch.equal(g_r[2], g_s[2])
    # the saved plans are what the backward used: with every tile's weights zeroed (headers kept) the
    # planned tiles contribute nothing to grad_value
    if sigma > 1:
        return
    tiles = 2 * 8 * ((421 + 63) // 64)
    stride, hdr = 14976 // 4, 640 // 4
    base = (nq * 4 + 255) // 256 * 64
    broken = plan_s.clone()
    broken.msda_saved = True
    broken[base:base + tiles * stride].view(tiles, stride)[:, hdr:] = 0
    g_b = MSDA.ms_deform_attn_backward(v, s, l, loc, aw, go, 64, plan=broken)
    assert float(g_b[0].float().abs().sum()) < 0.9 * float(g_s[0].float().abs().sum())


def test_prezero_on_and_off_give_the_same_gradients():
    """MSDeformAttnFunction with grad_value cleared ahead of the backward (PREZERO_GRAD_VALUE, default) and
    with the clear inside the backward: grad_sampling_loc / grad_attn_weight bitwise, grad_value up to the
    order of its atomics."""
    from mfl_b200.functions import ms_deform_attn_func as F
    inp = synthetic.make_inputs(2, 421, 8, 64, 4, [(1, 256), (1, 128), (1, 64), (1, 32)], torch.float32, DEV, seed=41,
                                dist="model", value="normal")
    grads = []
    for pre in (True, False):
        keep = F.PREZERO_GRAD_VALUE
        F.PREZERO_GRAD_VALUE = pre
        try:
            v, loc, aw = (inp[k].detach().clone().requires_grad_(True) for k in ("value", "loc", "aw"))
            F.ms_deform_attn(v, inp["shapes"], inp["lsi"], loc, aw, 64).backward(inp["grad_out"])
            grads.append((v.grad, loc.grad, aw.grad))
        finally:
            F.PREZERO_GRAD_VALUE = keep
    for k in (1, 2):
        dlt = (grads[0][k] - grads[1][k]).abs()
        assert torch.equal(grads[0][k], grads[1][k]), (k, float(dlt.max()), int((dlt > 0).sum()), dlt.numel())
    assert rel_err(grads[0][0].cpu().numpy(), grads[1][0].cpu().numpy()) <= 1e-5
