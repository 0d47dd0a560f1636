"""configs[4] harness == the reference's DeformableTransformer, parameter for parameter: the
state-dict keys and shapes of models/deformable/unimodal_deformable_transformer.py:13-60 (fixture
tests/golden/ref_transformer_keys.json, written by tests/golden/make_golden.py from the reference
itself) are exactly those of tools/train_step.py's DeformableTransformer; when /root/reference is
present (the build container) the reference model is instantiated and its state dict loaded."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))


def test_reference_transformer_state_dict_loads_key_for_key():
    import train_step
    mine = train_step.DeformableTransformer(512, 2048, 8, 4, 4, 6)
    want = json.load(open(os.path.join(ROOT, "tests", "golden", "ref_transformer_keys.json")))
    got = {k: list(v.shape) for k, v in mine.state_dict().items()}
    assert got == want
    assert sum(v.numel() for v in mine.state_dict().values()) == 39413249
    ref_dir = "/root/reference/models"
    if os.path.isdir(ref_dir):
        import importlib
        import types
        stub = types.ModuleType("models")
        stub.__path__ = [ref_dir]
        sys.modules["models"] = stub
        try:
            udt = importlib.import_module("models.deformable.unimodal_deformable_transformer")
            ref = udt.DeformableTransformer(d_model=512, num_head=8, num_encoder_layers=6, num_decoder_layers=6,
                                            dim_feedforward=2048, dropout=0.1, activation="relu",
                                            return_intermediate_dec=True, num_feature_levels=4, dec_n_points=4,
                                            enc_n_points=4)
            missing, unexpected = mine.load_state_dict(ref.state_dict(), strict=True)
            assert not missing and not unexpected
            k = "encoder.layers.3.self_attn.sampling_offsets.bias"
            assert torch.equal(mine.state_dict()[k], ref.state_dict()[k])
        finally:
            sys.modules.pop("models", None)
