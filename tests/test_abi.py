"""CPU suite: the C-ABI library builds, loads and exports every symbol
include/msda_b200.h declares; argument validation that needs no GPU."""
import ctypes
import os
import re

import pytest
import torch

import mfl_b200
from mfl_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    if not os.path.exists(_lib.LIB_PATH):
        _lib.build()
    return _lib.load()


def test_header_symbols_exported(lib):
    header = open(os.path.join(ROOT, "include", "msda_b200.h")).read()
    declared = set(re.findall(r"\b(msda_[a-z_0-9]+)\s*\(", header))
    assert declared == set(_lib.EXPORTED_SYMBOLS)
    for name in declared:
        assert hasattr(lib, name), name


def test_abi_version_and_strings(lib):
    assert lib.msda_abi_version() == _lib.ABI_VERSION == 2
    assert lib.msda_error_string(0) == b"ok"
    assert b"dtype" in lib.msda_error_string(-3)
    assert lib.msda_kernel_family(64, _lib.MSDA_F32, 0) == b"fast"
    assert lib.msda_kernel_family(64, _lib.MSDA_BF16, 0) == b"fast"
    assert lib.msda_kernel_family(64, _lib.MSDA_F64, 0) == b"generic"
    assert lib.msda_kernel_family(30, _lib.MSDA_F32, 0) == b"generic"
    assert lib.msda_kernel_family(64, _lib.MSDA_F32, _lib.MSDA_FLAG_FORCE_GENERIC) == b"generic"
    enc = (16, 1500, 8, 64, 4, 1500, 4)
    dec = (64, 1500, 8, 64, 4, 100, 4)
    assert lib.msda_kernel_choice(*enc, _lib.MSDA_F32, 0) == b"tile"
    assert lib.msda_kernel_choice(*enc, _lib.MSDA_BF16, 0) == b"tile"
    assert lib.msda_kernel_choice(*enc, _lib.MSDA_F32, _lib.MSDA_FLAG_NO_TILE) == b"window"
    assert lib.msda_forward_workspace_bytes(*enc, _lib.MSDA_F32, 0) == 16 * 1500 * 4
    tiles = 16 * 8 * ((1500 + 63) // 64)     # the plan of the split path: 13 KB per tile of 64 queries
    assert lib.msda_forward_workspace_bytes(*enc, _lib.MSDA_F32, _lib.MSDA_FLAG_PLAN) >= 16 * 1500 * 4 + tiles * 12288
    assert lib.msda_forward_workspace_bytes(*enc, _lib.MSDA_F32, _lib.MSDA_FLAG_NO_TILE) == 0
    assert lib.msda_backward_workspace_bytes_flags(*enc, _lib.MSDA_F32, 0) == 16 * 1500 * 4
    assert lib.msda_kernel_choice(*enc, _lib.MSDA_F32, _lib.MSDA_FLAG_NO_WINDOW) == b"fast"
    assert lib.msda_kernel_choice(*dec, _lib.MSDA_F32, 0) == b"fast"           # short query list
    assert lib.msda_kernel_choice(*dec, _lib.MSDA_F32, _lib.MSDA_FLAG_FORCE_WINDOW) == b"tile"
    assert lib.msda_forward_workspace_bytes(*dec, _lib.MSDA_F32, 0) == 0
    assert lib.msda_kernel_choice(*enc, _lib.MSDA_F64, 0) == b"generic"
    assert lib.msda_kernel_choice(16, 1500, 8, 64, 4, 1500, 8, _lib.MSDA_F32, 0) == b"fast"   # P != 4
    assert lib.msda_kernel_choice(16, 1500, 8, 64, 4, 1500, 4, 7, 0) == b"invalid"


def test_argument_validation_without_gpu(lib):
    # rejected before anything touches CUDA
    z = ctypes.c_void_p(0)
    assert lib.msda_forward(z, z, z, z, z, z, 1, 1, 0, 4, 1, 1, 1, 0, 0, 0, z) == -2   # M = 0
    assert lib.msda_forward(z, z, z, z, z, z, 1, 1, 1, 4, 1, 1, 1, 7, 0, 0, z) == -3   # dtype
    assert lib.msda_forward(z, z, z, z, z, z, 1, 1, 1, 4, 1, 1, 1, 0, 5, 0, z) == -4   # padding
    assert lib.msda_forward(z, z, z, z, z, z, 1, 1, 1, 4, 1, 1, 1, 0, 0, 0, z) == -1   # null ptrs
    assert lib.msda_forward(z, z, z, z, z, z, 1, 2 ** 31, 1, 4, 1, 1, 1, 0, 0, 0, z) == -5
    assert lib.msda_forward(z, z, z, z, z, z, 0, 1, 1, 4, 1, 1, 1, 0, 0, 0, z) == 0    # empty batch
    assert lib.msda_backward_workspace_bytes(2, 3, 4, 8, 1, 1, 1, _lib.MSDA_F32) == 0
    assert lib.msda_backward_workspace_bytes(2, 3, 4, 8, 1, 1, 1, _lib.MSDA_BF16) == 2 * 3 * 4 * 8 * 4
    # msda_forward_static (no reference counterpart): the same validation, and a call whose plan
    # does not fit the built-in 1 MiB array is refused on the host, before any CUDA call
    assert lib.msda_forward_static(z, z, z, z, z, z, 1, 1, 0, 4, 1, 1, 1, 0, 0, 0, z) == -2
    assert lib.msda_forward_static(z, z, z, z, z, z, 0, 1, 1, 4, 1, 1, 1, 0, 0, 0, z) == 0
    big = (176, 1500, 8, 64, 4, 1500, 4)                      # 264,000 queries > 262,144
    assert lib.msda_forward_workspace_bytes(*big, _lib.MSDA_F32, 0) > (1 << 20)
    assert lib.msda_forward_static(z, z, z, z, z, z, *big, _lib.MSDA_F32, 0, 0, z) == -11
    assert b"scratch" in lib.msda_error_string(-11).lower()


def _cpu_inputs():
    value = torch.rand(2, 6, 2, 4)
    shapes = torch.tensor([[1, 4], [1, 2]])
    lsi = torch.tensor([0, 4])
    loc = torch.rand(2, 3, 2, 2, 2, 2)
    aw = torch.rand(2, 3, 2, 2, 2)
    return value, shapes, lsi, loc, aw


def test_cpu_tensors_raise_like_reference():
    # models/ops/src/ms_deform_attn.h:38 -- AT_ERROR("Not implemented on the CPU")
    with pytest.raises(RuntimeError, match="Not implemented on the CPU"):
        mfl_b200.MSDeformAttnFunction.apply(*_cpu_inputs(), 64)
    m = mfl_b200.MSDeformAttn(d_model=32, n_levels=2, n_heads=4, n_points=2)
    with pytest.raises(RuntimeError, match="Not implemented on the CPU"):
        m(torch.rand(1, 3, 32), torch.rand(1, 3, 2, 1), torch.rand(1, 6, 32),
          torch.tensor([4, 2]), torch.tensor([0, 4]))


def test_module_surface_matches_reference():
    # state-dict keys / shapes; ctor errors (ms_deform_attn.py:40-41)
    m = mfl_b200.MSDeformAttn(d_model=512, n_levels=4, n_heads=8, n_points=4)
    sd = {k: tuple(v.shape) for k, v in m.state_dict().items()}
    assert sd == {
        "sampling_offsets.weight": (128, 512), "sampling_offsets.bias": (128,),
        "attention_weights.weight": (128, 512), "attention_weights.bias": (128,),
        "value_proj.weight": (512, 512), "value_proj.bias": (512,),
        "output_proj.weight": (512, 512), "output_proj.bias": (512,)}
    assert m.im2col_step == 64
    with pytest.raises(ValueError):
        mfl_b200.MSDeformAttn(d_model=30, n_heads=8)
    # bias grid: head direction cos(2*pi*m/M) sign-normalised, times (p+1)  (ms_deform_attn.py:64-71)
    b = m.sampling_offsets.bias.detach().view(8, 4, 4)
    assert torch.allclose(b[0, 0], torch.tensor([1., 2., 3., 4.]))
    assert torch.allclose(b[4, 2], torch.tensor([-1., -2., -3., -4.]))
    assert float(b[2].abs().max()) < 1e-6


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "multimodal-feature-learning_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".sh")):
                text = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in text and "from oracle" not in text, f
                if f.endswith(".py"):
                    assert "F.grid_sample(" not in text, f


def test_reference_names_resolve():
    """models/ops/functions/__init__.py:9 and `import MultiScaleDeformableAttention as MSDA`
    (ms_deform_attn_func.py:18-21) resolve against this repo; CPU tensors raise."""
    import torch
    from mfl_b200.functions import MSDeformAttnFunction, ms_deform_attn_core_pytorch  # noqa: F401
    import MultiScaleDeformableAttention as MSDA
    assert callable(MSDA.ms_deform_attn_forward) and callable(MSDA.ms_deform_attn_backward)
    v = torch.zeros(1, 6, 2, 4)
    loc = torch.zeros(1, 3, 2, 1, 2, 2)
    aw = torch.zeros(1, 3, 2, 1, 2)
    with pytest.raises(RuntimeError, match="Not implemented on the CPU"):
        ms_deform_attn_core_pytorch(v, [(2, 3)], loc, aw)


def test_workspace_sizes_host_logic(lib):
    """The workspace queries are pure host arithmetic (include/msda_b200.h): the plan is one int per
    query, kernel families without a plan ask for nothing, the backward's required part (fp32 image
    for bf16, fixed-point image + header for the deterministic variant) comes first, 256-byte rounded,
    with the plan behind it -- and msda_forward_static's limit is exactly 262,144 queries."""
    enc = (16, 1500, 8, 64, 4, 1500, 4)
    dec = (64, 1500, 8, 64, 4, 100, 4)
    fwd = lib.msda_forward_workspace_bytes
    bwd = lib.msda_backward_workspace_bytes_flags
    assert fwd(*enc, _lib.MSDA_F32, 0) == fwd(*enc, _lib.MSDA_BF16, 0) == 16 * 1500 * 4
    assert fwd(*dec, _lib.MSDA_F32, 0) == 0 and fwd(*enc, _lib.MSDA_F64, 0) == 0      # per-tap / generic
    assert fwd(*enc, _lib.MSDA_F32, _lib.MSDA_FLAG_NO_TILE) == 0
    assert fwd(1, 1500, 8, 64, 4, 262144, 4, _lib.MSDA_F32, 0) == 1 << 20
    assert fwd(1, 1500, 8, 64, 4, 262145, 4, _lib.MSDA_F32, 0) > 1 << 20
    plan = 16 * 1500 * 4
    nvalue = 16 * 1500 * 8 * 64
    assert bwd(*enc, _lib.MSDA_F32, 0) == plan
    assert bwd(*enc, _lib.MSDA_BF16, 0) == nvalue * 4 + plan                          # already a 256 multiple
    det = bwd(*enc, _lib.MSDA_F32, _lib.MSDA_FLAG_DETERMINISTIC)
    assert det > nvalue * 8 + plan and (det - plan) % 256 == 0 and det - plan - nvalue * 8 < 512
    assert lib.msda_backward_workspace_bytes(*enc, _lib.MSDA_F32) == bwd(*enc, _lib.MSDA_F32, 0)
    for bad in ((16, 1500, 0, 64, 4, 1500, 4), (16, 1500, 8, 64, 4, 1500, 0)):         # invalid shapes ask for nothing
        assert fwd(*bad, _lib.MSDA_F32, 0) == 0
