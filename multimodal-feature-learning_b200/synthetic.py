"""Seeded synthetic inputs for the op (tests and bench.py) -- SURVEY.md section 8(d).

Workloads are named after BASELINE.json's configs:
  CPU : configs[0]  N=2,  Lq=100,  T=[300,150,75,38]   (the reference's CPU-runnable case)
  ENC : configs[1]  N=16, Lq=S=1500, T=[800,400,200,100] (encoder self-attention)
  DEC : configs[3]  N=64, Lq=100,  same T              (decoder cross-attention)
all with M=8 heads, D=64 channels/head, L=4 temporal levels, P=4 points.
"""
import math

import torch

WORKLOADS = {
    "CPU": dict(N=2, Lq=100, T=[300, 150, 75, 38], M=8, D=64, P=4, enc=False),
    "ENC": dict(N=16, Lq=None, T=[800, 400, 200, 100], M=8, D=64, P=4, enc=True),
    "DEC": dict(N=64, Lq=100, T=[800, 400, 200, 100], M=8, D=64, P=4, enc=False),
    # the four calls of one multimodal encoder layer
    # (models/deformable/multimodal_deformable_transformer.py:256-271): video and audio
    # self-attention, then audio tokens over the video memory and video tokens over the audio memory
    "MM_VV": dict(N=16, Lq=None, T=[640, 320, 160, 80], M=8, D=64, P=4, enc=True),
    "MM_AA": dict(N=16, Lq=None, T=[160, 80, 40, 20], M=8, D=64, P=4, enc=True),
    "MM_AV": dict(N=16, Lq=None, T=[640, 320, 160, 80], M=8, D=64, P=4, enc=True, Tq=[160, 80, 40, 20]),
    "MM_VA": dict(N=16, Lq=None, T=[160, 80, 40, 20], M=8, D=64, P=4, enc=True, Tq=[640, 320, 160, 80]),
}


def level_tensors(shapes_hw, device):
    shapes = torch.as_tensor(shapes_hw, dtype=torch.long, device=device).reshape(-1, 2)
    sizes = shapes.prod(1)
    lsi = torch.cat((sizes.new_zeros((1,)), sizes.cumsum(0)[:-1]))
    return shapes, lsi, int(sizes.sum())


def head_bias(M, P, device):
    """sampling-offset bias of a freshly initialised module
    (models/ops/modules/ms_deform_attn.py:64-71): sign-normalised cos(2*pi*m/M) * (p+1)."""
    ang = torch.arange(M, dtype=torch.float32, device=device) * (2.0 * math.pi / M)
    ring = torch.stack([ang.cos(), ang.sin()], -1)
    x = (ring / ring.abs().max(-1, keepdim=True)[0])[:, 0]
    return x[:, None] * torch.arange(1, P + 1, dtype=torch.float32, device=device)[None, :]


def encoder_reference_points(T, device):
    """(t + 0.5) / T_l for every token of every level (valid_ratio = 1)
    -- models/deformable/unimodal_deformable_transformer.py:259-276."""
    return torch.cat([(torch.arange(t, dtype=torch.float32, device=device) + 0.5) / t for t in T])


def make_inputs(N, Lq, M, D, P, shapes_hw, dtype=torch.float32, device="cuda", seed=1234,
                dist="uniform", value="small", enc_ref=False, requires_grad=False, query_T=None,
                sigma=0.5, permute=False):
    """dist: 'uniform'  loc ~ U[0,1)^2 (models/ops/test.py:34)
             'model'    loc_x = ref + (bias[m,p] + 0.5*N(0,1)) / W_l   (a freshly initialised module)
             'interior' every tap inside the map (zeros == border)
             'spill'    loc ~ U[-0.25, 1.25) (exercises the gate / clamping)
       value: 'small' ~ U[0,1)*0.01 (test.py:33), 'normal' ~ N(0,1)
       enc_ref: the queries are the tokens of a multi-level sequence (of the value's own
       levels, or of `query_T` when the queries come from the other modality) and carry the
       encoder's reference points.
       sigma: rows of Gaussian noise on the 'model' offsets (0.5 = a freshly initialised module;
       larger values probe how much learned offset spread the windowed kernels take).
       permute: shuffle the query list per batch item (the sparse encoders hand the op
       score-ordered top-k tokens: models/sparse/unimodal_sparse_deformable_transformer.py:218).
       For maps with H == 1 the y coordinate is the constant 0.5 the module stacks on."""
    gen = torch.Generator(device=device).manual_seed(seed)
    shapes, lsi, S = level_tensors(shapes_hw, device)
    L = shapes.shape[0]
    f32 = dict(dtype=torch.float32, device=device, generator=gen)
    if value == "small":
        val = torch.rand(N, S, M, D, **f32) * 0.01
    else:
        val = torch.randn(N, S, M, D, **f32)
    if enc_ref:
        ref = encoder_reference_points(query_T or [int(w) for _, w in shapes_hw], device)[None, :].expand(N, -1)
        assert ref.shape[1] == Lq
    else:
        ref = torch.rand(N, Lq, **f32)
    Wl = shapes[:, 1].float()[None, None, None, :, None]
    Hl = shapes[:, 0].float()[None, None, None, :, None]
    if dist == "model":
        off = head_bias(M, P, device)[None, None, :, None, :] + sigma * torch.randn(N, Lq, M, L, P, **f32)
        x = ref[:, :, None, None, None] + off / Wl
        y = torch.rand(N, Lq, M, L, P, **f32)
    else:
        x = torch.rand(N, Lq, M, L, P, **f32)
        y = torch.rand(N, Lq, M, L, P, **f32)
        if dist == "interior":
            x = (0.5 + x * (Wl - 1)) / Wl
            y = (0.5 + y * (Hl - 1)) / Hl
        elif dist == "spill":
            x = x * 1.5 - 0.25
            y = y * 1.5 - 0.25
        elif dist != "uniform":
            raise ValueError(dist)
    y = torch.where(Hl.expand_as(y) == 1, torch.full_like(y, 0.5), y)
    loc = torch.stack((x, y), -1)
    aw = torch.softmax(torch.randn(N, Lq, M, L * P, **f32), -1).view(N, Lq, M, L, P)
    gout = torch.randn(N, Lq, M * D, **f32)
    if permute:
        idx = torch.stack([torch.randperm(Lq, device=device, generator=gen) for _ in range(N)])
        loc = torch.gather(loc, 1, idx[:, :, None, None, None, None].expand_as(loc))
        aw = torch.gather(aw, 1, idx[:, :, None, None, None].expand_as(aw))
    out = dict(value=val.to(dtype), shapes=shapes, lsi=lsi, loc=loc.to(dtype).contiguous(),
               aw=aw.to(dtype).contiguous(), grad_out=gout.to(dtype))
    if requires_grad:
        for k in ("value", "loc", "aw"):
            out[k].requires_grad_(True)
    return out


def make_workload(name, dtype=torch.float32, device="cuda", seed=1234, dist="model", value="normal",
                  N=None, sigma=0.5, permute=False):
    w = WORKLOADS[name]
    T = w["T"]
    Lq = w["Lq"] if w["Lq"] is not None else sum(w.get("Tq", T))
    return make_inputs(N or w["N"], Lq, w["M"], w["D"], w["P"], [(1, t) for t in T], dtype, device,
                       seed, dist, value, enc_ref=w["enc"], query_T=w.get("Tq"), sigma=sigma, permute=permute)


def algorithmic_bytes(N, S, Lq, M, D, L, P, elem_bytes):
    """the single byte formula for roofline.achieved."""
    fwd = elem_bytes * (N * S * M * D + 3 * N * Lq * M * L * P + N * Lq * M * D)
    bwd = elem_bytes * (N * Lq * M * D + N * S * M * D + 3 * N * Lq * M * L * P
                        + N * S * M * D + 3 * N * Lq * M * L * P)
    return fwd, bwd
