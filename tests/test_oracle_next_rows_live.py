"""CPU suite, build container only: the oracles of the "next" rows (decoder attention map,
encoder reference points) against the reference's own functions imported unmodified from
/root/reference, on seeded random shapes beyond the one golden fixture each has
(tests/golden/dam_f32.npz, refpoints_f32.npz).  Skips on the GPU box, which has no reference."""
import importlib
import importlib.util
import os
import sys
import types

import pytest
import torch

from oracle import core_pytorch, dam_pytorch

REF = "/root/reference"
pytestmark = pytest.mark.skipif(not os.path.isdir(REF), reason="/root/reference is only in the build container")


def _load(path, name):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


@pytest.mark.parametrize("seed,B,NL,Q,M,T,P", [
    (1, 1, 1, 1, 1, [5], 1),
    (2, 2, 6, 100, 8, [800, 400, 200, 100][::1], 4),
    (3, 3, 2, 7, 3, [9, 1, 4], 2),
    (4, 1, 2, 13, 2, [31, 16], 5),
])
def test_dam_restatement_equals_reference_dam(seed, B, NL, Q, M, T, P):
    """utils/dam.py:20-73, locations spilling past both ends of the levels."""
    ref_dam = _load(os.path.join(REF, "utils/dam.py"), "_ref_utils_dam_live")
    torch.manual_seed(seed)
    T = torch.tensor(T, dtype=torch.long)
    L = len(T)
    lsi = torch.cat((T.new_zeros(1), T.cumsum(0)[:-1]))
    loc = torch.rand(B, NL, Q, M, L, P, 1) * 1.4 - 0.2
    aw = torch.softmax(torch.randn(B, NL, Q, M, L * P), -1).view(B, NL, Q, M, L, P)
    want = ref_dam.attn_map_to_flat_grid(T, lsi, loc, aw)
    got = dam_pytorch.attn_map_to_flat_grid(T, lsi, loc, aw)
    assert got.shape == want.shape
    # scatter_add order may differ between the two expressions: fp32 rounding only
    torch.testing.assert_close(got, want, rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("seed,N,T", [(1, 1, [7]), (2, 3, [24, 12, 6, 3]), (3, 16, [800, 400, 200, 100]),
                                      (4, 2, [1, 5, 1])])
def test_reference_points_restatement_equals_reference_method(seed, N, T):
    """DeformableTransformerEncoder.get_reference_points
    (models/deformable/unimodal_deformable_transformer.py:259-276), imported through a stub `models`
    package as tests/golden/make_golden.py does (models/__init__.py needs packages this image lacks)."""
    udt = _ref_udt()
    torch.manual_seed(seed)
    T = torch.tensor(T, dtype=torch.long)
    vr = torch.rand(N, len(T)) * 0.6 + 0.4
    want = udt.DeformableTransformerEncoder.get_reference_points(T, vr, "cpu")
    got = core_pytorch.reference_points_pytorch(T, vr)
    assert got.shape == want.shape
    torch.testing.assert_close(got, want, rtol=1e-6, atol=1e-7)


def _ref_udt():
    saved = sys.modules.get("models")
    stub = types.ModuleType("models")
    stub.__path__ = [os.path.join(REF, "models")]
    sys.modules["models"] = stub
    try:
        return importlib.import_module("models.deformable.unimodal_deformable_transformer")
    finally:
        for k in [k for k in sys.modules if k == "models" or k.startswith("models.")]:
            del sys.modules[k]
        if saved is not None:
            sys.modules["models"] = saved


@pytest.mark.parametrize("sizes,C,B", [([300, 150, 75, 38], 64, 2), ([33, 7], 48, 3), ([5], 8, 1)])
def test_prepare_encoder_inputs_checker_equals_reference_method(sizes, C, B):
    """The checker the GPU test of msda_prepare_encoder_inputs compares with (ref_prepare in
    tests/test_prepare_inputs_gpu.py) against DeformableTransformer.prepare_encoder_inputs itself
    (models/deformable/unimodal_deformable_transformer.py:90-134), called unbound on a stand-in that
    carries only what the method reads (level_embed, get_valid_ratio): all six results bit for bit."""
    from test_prepare_inputs_gpu import ref_prepare
    udt = _ref_udt()
    cls = udt.DeformableTransformer
    g = torch.Generator().manual_seed(5)
    srcs = [torch.randn(B, C, t, generator=g) for t in sizes]
    poss = [torch.randn(B, C, t, generator=g) for t in sizes]
    masks = [torch.rand(B, t, generator=g) > 0.8 for t in sizes]
    le = torch.randn(len(sizes), C, generator=g)
    me = types.SimpleNamespace(level_embed=le)
    me.get_valid_ratio = lambda m: cls.get_valid_ratio(me, m)
    want = cls.prepare_encoder_inputs(me, srcs, masks, poss)
    got = ref_prepare(srcs, masks, poss, le)
    assert len(got) == len(want) == 6
    for a, b in zip(got, want):
        assert a.shape == b.shape and a.dtype == b.dtype and torch.equal(a, b)
