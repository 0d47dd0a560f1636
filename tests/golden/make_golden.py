"""Generate the golden fixtures in tests/golden/*.npz FROM THE REFERENCE ITSELF.

Run in the build container only (needs /root/reference, which does not exist
on the GPU box):      python tests/golden/make_golden.py

It imports, by file path and unmodified,
  * models/ops/functions/ms_deform_attn_func.py  (ms_deform_attn_core_pytorch, 2-D form, :44-71)
  * models/modules/attention.py                  (ms_deform_attn_core_pytorch, 1-D live twin, :331-383;
                                                   MSDeformAttn, the module every transformer builds, :394-511)
  * utils/dam.py                                 (attn_map_to_flat_grid, :20-73)
  * models/deformable/unimodal_deformable_transformer.py (get_reference_points, :259-276)
runs them on seeded inputs (the tiny case reproduces models/ops/test.py:23-44:
N=1, M=2, D=2, Lq=2, L=2, P=2, shapes (6,4),(3,2), torch.manual_seed(3)) and
stores inputs, outputs and autograd gradients.

Two padding variants are stored per case:
  border : the reference function exactly as written (padding_mode='border').
  zeros  : the SAME reference function executed with F.grid_sample's
           padding_mode argument overridden to 'zeros' -- the CPU-available
           stand-in for the reference CUDA kernel's zero padding
           (ms_deform_im2col_cuda.cuh:56-79, :289); SURVEY.md section 8(c) oracle (3).
"""
import importlib.util
import os
import sys

import numpy as np
import torch
import torch.nn.functional as F

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))


def _load(path, name):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


ref_ops = _load(os.path.join(REF, "models/ops/functions/ms_deform_attn_func.py"), "ref_ops_func")
ref_live = _load(os.path.join(REF, "models/modules/attention.py"), "ref_live_attention")

_real_grid_sample = F.grid_sample


class _force_padding:
    """Run the reference with grid_sample(padding_mode=<mode>) -- one argument differs."""

    def __init__(self, mode):
        self.mode = mode

    def __enter__(self):
        def patched(inp, grid, mode="bilinear", padding_mode="zeros", align_corners=None):
            return _real_grid_sample(inp, grid, mode=mode, padding_mode=self.mode,
                                     align_corners=align_corners)
        F.grid_sample = patched

    def __exit__(self, *a):
        F.grid_sample = _real_grid_sample


def run_ref(fn, value, shapes_arg, loc_arg, aw, grad_out, padding):
    value = value.clone().requires_grad_(True)
    loc_arg = loc_arg.clone().requires_grad_(True)
    aw = aw.clone().requires_grad_(True)
    if padding == "border":
        out = fn(value, shapes_arg, loc_arg, aw)
    else:
        with _force_padding(padding):
            out = fn(value, shapes_arg, loc_arg, aw)
    out.backward(grad_out)
    return out.detach(), value.grad, loc_arg.grad, aw.grad


def make_case(name, seed, N, M, D, Lq, P, shapes_hw, dtype, loc_range=(0.0, 1.0), interior=False,
              one_d_twin=False, value_scale=0.01):
    torch.manual_seed(seed)
    shapes = torch.as_tensor(shapes_hw, dtype=torch.long)
    L = shapes.shape[0]
    lsi = torch.cat((shapes.new_zeros((1,)), shapes.prod(1).cumsum(0)[:-1]))
    S = int(shapes.prod(1).sum())
    value = (torch.rand(N, S, M, D) * value_scale).to(dtype)
    lo, hi = loc_range
    loc = (torch.rand(N, Lq, M, L, P, 2) * (hi - lo) + lo).to(dtype)
    if interior:  # keep every tap inside the map so zeros == border
        wh = torch.stack((shapes[:, 1], shapes[:, 0]), -1).to(dtype)[None, None, None, :, None, :]
        loc = (0.5 + loc * (wh - 1)) / wh
    if one_d_twin:
        assert all(h == 1 for h, _ in shapes_hw)
        loc[..., 1] = 0.5
    aw = torch.rand(N, Lq, M, L, P) + 1e-5
    aw = (aw / aw.sum(-1, keepdim=True).sum(-2, keepdim=True)).to(dtype)
    grad_out = torch.randn(N, Lq, M * D).to(dtype)
    rec = dict(value=value, shapes=shapes, lsi=lsi, loc=loc, aw=aw, grad_out=grad_out)
    for padding in ("border", "zeros"):
        o, gv, gl, ga = run_ref(ref_ops.ms_deform_attn_core_pytorch, value, shapes, loc, aw, grad_out,
                                padding)
        rec.update({"out_" + padding: o, "gvalue_" + padding: gv, "gloc_" + padding: gl,
                    "gaw_" + padding: ga})
    if one_d_twin:
        # the live 1-D twin takes shapes [L,1] (T_l) and loc [...,1]; must agree with the 2-D form
        o, gv, gl, ga = run_ref(ref_live.ms_deform_attn_core_pytorch, value, shapes[:, 1:].contiguous(),
                                loc[..., :1].contiguous(), aw, grad_out, "border")
        rec.update(dict(out_live=o, gvalue_live=gv, gloc_live=gl, gaw_live=ga))
    np.savez_compressed(os.path.join(HERE, name + ".npz"),
                        **{k: v.numpy() for k, v in rec.items()})
    print(name, {k: tuple(v.shape) for k, v in rec.items() if k.startswith(("value", "loc"))})


def make_module_case(name, seed, rd, with_mask, encoder=False):
    """The reference's live MSDeformAttn (attention.py:394-511) run on CPU: parameters, inputs,
    output, the (loc, aw) it returns with is_sparse=True and the gradients of sum(out * g).
    encoder=True: an encoder-class call (one 64-channel head, 270 tokens attending to themselves
    with the encoder's reference points) -- the shape class of the fused module op."""
    torch.manual_seed(seed)
    C, L, M, P = (64, 4, 1, 4) if encoder else (64, 4, 4, 4)
    T = torch.tensor([144, 72, 36, 18] if encoder else [24, 12, 6, 3], dtype=torch.long)
    lsi = torch.cat((T.new_zeros(1), T.cumsum(0)[:-1]))
    N, Lq, S = (1, int(T.sum()), int(T.sum())) if encoder else (2, 17, int(T.sum()))
    mod = ref_live.MSDeformAttn(C, L, M, P)
    with torch.no_grad():
        mod.sampling_offsets.weight.normal_(0, 0.05)     # otherwise the offsets do not depend on the query
    q = torch.randn(N, Lq, C, requires_grad=True)
    src = torch.randn(N, S, C, requires_grad=True)
    if encoder:
        ref = torch.cat([(torch.arange(int(t)) + 0.5) / int(t) for t in T])[None, :, None, None].expand(N, -1, L, 1).clone()
    else:
        ref = torch.rand(N, Lq, L, rd)
    if rd == 2:
        ref[..., 1] = ref[..., 1] * 0.5 + 0.05
    ref.requires_grad_(True)
    mask = None
    if with_mask:
        mask = torch.zeros(N, S, dtype=torch.bool)
        mask[N - 1, -9:] = True
    out, loc, aw = mod(q, ref, src, T, lsi, mask, is_sparse=True)
    g = torch.randn_like(out)
    (out * g).sum().backward()
    rec = {"param." + k: v.detach() for k, v in mod.state_dict().items()}
    rec.update(query=q.detach(), src=src.detach(), ref=ref.detach(), T=T, lsi=lsi, g=g, out=out.detach(),
               loc=loc.detach(), aw=aw.detach(), g_query=q.grad, g_src=src.grad, g_ref=ref.grad,
               mask=mask if mask is not None else torch.zeros(0, dtype=torch.bool))
    rec.update({"gparam." + k: p.grad for k, p in mod.named_parameters()})
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **{k: v.numpy() for k, v in rec.items()})
    print(name, tuple(out.shape), tuple(loc.shape))


def make_dam_case(name, seed):
    """utils/dam.py:20-73 on seeded decoder-like inputs, locations spilling past both ends."""
    ref_dam = _load(os.path.join(REF, "utils/dam.py"), "ref_utils_dam")
    torch.manual_seed(seed)
    B, NL, Q, M, L, P = 2, 3, 10, 4, 4, 4
    T = torch.tensor([24, 12, 6, 3], dtype=torch.long)
    lsi = torch.cat((T.new_zeros(1), T.cumsum(0)[:-1]))
    loc = torch.rand(B, NL, Q, M, L, P, 1) * 1.3 - 0.15
    aw = torch.softmax(torch.randn(B, NL, Q, M, L * P), -1).view(B, NL, Q, M, L, P)
    out = ref_dam.attn_map_to_flat_grid(T, lsi, loc, aw)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), T=T.numpy(), lsi=lsi.numpy(), loc=loc.numpy(),
                        aw=aw.numpy(), out=out.numpy())
    print(name, tuple(out.shape))


def make_refpoints_case(name, seed):
    """DeformableTransformerEncoder.get_reference_points
    (models/deformable/unimodal_deformable_transformer.py:259-276), imported through a stub
    `models` package (models/__init__.py pulls in dependencies this container lacks)."""
    import importlib
    import types
    stub = types.ModuleType("models")
    stub.__path__ = [os.path.join(REF, "models")]
    sys.modules["models"] = stub
    udt = importlib.import_module("models.deformable.unimodal_deformable_transformer")
    torch.manual_seed(seed)
    T = torch.tensor([24, 12, 6, 3], dtype=torch.long)
    vr = torch.rand(3, 4) * 0.6 + 0.4
    out = udt.DeformableTransformerEncoder.get_reference_points(T, vr, "cpu")
    np.savez_compressed(os.path.join(HERE, name + ".npz"), T=T.numpy(), vr=vr.numpy(), out=out.numpy())
    print(name, tuple(out.shape))


def main():
    make_refpoints_case("refpoints_f32", 61)
    make_module_case("module_live_ref1_f32", 41, 1, True)
    make_module_case("module_live_ref2_f32", 42, 2, False)
    make_module_case("module_live_enc_f32", 43, 1, True, encoder=True)
    make_dam_case("dam_f32", 51)
    # models/ops/test.py:17-44 (the only shapes/seeds the reference's tests fix)
    make_case("testpy_tiny_f64", 3, 1, 2, 2, 2, 2, [(6, 4), (3, 2)], torch.float64)
    make_case("testpy_tiny_f32", 3, 1, 2, 2, 2, 2, [(6, 4), (3, 2)], torch.float32)
    # 2-D, odd channel count, locations spilling outside the map (gate / clamping edge cases)
    make_case("edge2d_f64", 11, 2, 3, 5, 7, 3, [(5, 7), (3, 4), (1, 2)], torch.float64,
              loc_range=(-0.3, 1.3))
    # temporal (H == 1) use of the op, as every transformer in the reference calls it
    make_case("temporal_f64", 21, 2, 4, 8, 12, 4, [(1, 16), (1, 8), (1, 4), (1, 2)], torch.float64,
              loc_range=(-0.1, 1.1), one_d_twin=True)
    make_case("temporal_f32", 22, 2, 8, 16, 20, 4, [(1, 38), (1, 19), (1, 10), (1, 5)], torch.float32,
              loc_range=(-0.1, 1.1), one_d_twin=True)
    # interior-only locations: zeros and border coincide
    make_case("interior2d_f32", 31, 2, 2, 8, 9, 4, [(6, 5), (3, 3)], torch.float32, interior=True)


if __name__ == "__main__":
    if not os.path.isdir(REF):
        sys.exit("needs %s (build container only)" % REF)
    main()


def make_transformer_keys():
    """State-dict keys and shapes of the reference's DeformableTransformer at the configs[4] size
    (models/deformable/unimodal_deformable_transformer.py:13-60) -> ref_transformer_keys.json;
    tools/train_step.py's model must accept them key for key."""
    import importlib
    import json
    import types
    stub = types.ModuleType("models")
    stub.__path__ = [os.path.join(REF, "models")]
    sys.modules["models"] = stub
    udt = importlib.import_module("models.deformable.unimodal_deformable_transformer")
    m = udt.DeformableTransformer(d_model=512, num_head=8, num_encoder_layers=6, num_decoder_layers=6,
                                  dim_feedforward=2048, dropout=0.1, activation="relu", return_intermediate_dec=True,
                                  num_feature_levels=4, dec_n_points=4, enc_n_points=4)
    with open(os.path.join(HERE, "ref_transformer_keys.json"), "w") as f:
        json.dump({k: list(v.shape) for k, v in m.state_dict().items()}, f, indent=0)


if os.environ.get("MSDA_GOLDEN_TRANSFORMER_KEYS"):
    make_transformer_keys()
