"""msda_forward_static: the forward for callers that cannot allocate (include/msda_b200.h).

No counterpart in the reference.  The plan of the
CTA-tile kernels lives in a file-scope __device__ array of MSDA_STATIC_SCRATCH_BYTES inside
libmsda_b200.so instead of a caller workspace.  Checked here through the C ABI against the
existing entry point (msda_forward_ex with a workspace, via the Python shim), on two calls in a
row with different shapes and inputs (the array keeps the first call's plan: the second call
must not see it), plus the size limit and a shape class that needs no scratch.

Collected last on purpose (file name).
Tolerance as in tests/test_parity_gpu.py / tests/test_window_gpu.py.
"""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

import mfl_b200  # noqa: E402,F401
from mfl_b200 import MultiScaleDeformableAttention as MSDA  # noqa: E402
from mfl_b200 import _lib, synthetic  # noqa: E402

DEV = "cuda"
FWD_ABS_F32 = 1e-5
BF16_FWD_REL = 2.0 ** -7
STATIC_SCRATCH_BYTES = 1 << 20          # MSDA_STATIC_SCRATCH_BYTES
ERR_SCRATCH_LIMIT = -11                 # MSDA_ERR_SCRATCH_LIMIT
DTYPES = {torch.float32: _lib.MSDA_F32, torch.bfloat16: _lib.MSDA_BF16}


def dims_of(inp):
    N, S, M, D = inp["value"].shape
    Lq, L, P = inp["loc"].shape[1], inp["loc"].shape[3], inp["loc"].shape[4]
    return N, S, M, D, L, Lq, P


def forward_static(inp, padding="zeros", flags=0):
    """The new entry point, raw: the output buffer is the caller's, the call allocates nothing."""
    lib = _lib.load()
    N, S, M, D, L, Lq, P = dims_of(inp)
    out = torch.full((N, Lq, M * D), float("nan"), dtype=inp["value"].dtype, device=DEV)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = lib.msda_forward_static(inp["value"].data_ptr(), inp["shapes"].data_ptr(), inp["lsi"].data_ptr(),
                                 inp["loc"].data_ptr(), inp["aw"].data_ptr(), out.data_ptr(),
                                 N, S, M, D, L, Lq, P, DTYPES[inp["value"].dtype],
                                 _lib.PADDING_MODES[padding], flags, stream)
    assert rc == 0, lib.msda_error_string(rc).decode()
    return out


def forward_existing(inp, padding="zeros", flags=0):
    return MSDA.ms_deform_attn_forward(inp["value"], inp["shapes"], inp["lsi"], inp["loc"], inp["aw"], 64,
                                       padding_mode=padding, flags=flags)


def choice(inp, flags=0):
    lib = _lib.load()
    return lib.msda_kernel_choice(*dims_of(inp), DTYPES[inp["value"].dtype], flags).decode()


def assert_close(a, b, dtype):
    a = a.float().cpu().numpy()
    b = b.float().cpu().numpy()
    assert np.isfinite(a).all()
    if dtype == torch.float32:
        assert np.abs(a - b).max() <= FWD_ABS_F32
    else:
        assert np.abs(a - b).max() <= BF16_FWD_REL * max(np.abs(b).max(), 1e-30)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_two_calls_in_a_row_match_the_workspace_entry_point(dtype):
    # first call: 3 batch items of 700 queries; second: 2 items of 1203 (ragged last tile), other
    # seed, other padding -- the array still holds the first call's plan when the second starts
    a = synthetic.make_inputs(3, 700, 8, 64, 4, [(1, 400), (1, 200)], dtype=dtype, device=DEV, seed=11,
                              dist="model", value="normal")
    b = synthetic.make_inputs(2, 1203, 8, 64, 4, [(1, 640), (1, 320), (1, 160), (1, 83)], dtype=dtype, device=DEV,
                              seed=12, dist="model", value="normal", permute=True)
    assert choice(a) == "tile" and choice(b) == "tile"      # the shape class that uses scratch
    want_a = forward_existing(a, "zeros")
    want_b = forward_existing(b, "border")
    got_a = forward_static(a, "zeros")
    got_b = forward_static(b, "border")
    got_a2 = forward_static(a, "zeros")                      # and back again
    torch.cuda.synchronize()
    assert_close(got_a, want_a, dtype)
    assert_close(got_b, want_b, dtype)
    assert_close(got_a2, want_a, dtype)


def test_plan_flag_is_ignored_and_no_scratch_shapes_run():
    a = synthetic.make_inputs(2, 512, 8, 32, 4, [(1, 256), (1, 128)], device=DEV, seed=13, dist="model",
                              value="normal")
    got = forward_static(a, "zeros", flags=_lib.MSDA_FLAG_PLAN)
    want = forward_existing(a, "zeros")
    # decoder-class call: per-tap kernels, no scratch at all
    d = synthetic.make_inputs(4, 100, 8, 64, 4, [(1, 300), (1, 150), (1, 75), (1, 38)], device=DEV, seed=14,
                              dist="uniform")
    assert choice(d) != "tile"
    got_d = forward_static(d, "zeros")
    want_d = forward_existing(d, "zeros")
    torch.cuda.synchronize()
    assert_close(got, want, torch.float32)
    assert_close(got_d, want_d, torch.float32)


def test_size_limit_is_an_error_and_enqueues_nothing():
    lib = _lib.load()
    N, Lq = 64, 4097                      # 262,208 queries: 256 B over the array
    S, M, D, L, P = 1500, 8, 64, 4, 4
    need = lib.msda_forward_workspace_bytes(N, S, M, D, L, Lq, P, _lib.MSDA_F32, 0)
    assert need > STATIC_SCRATCH_BYTES
    rc = lib.msda_forward_static(None, None, None, None, None, None, N, S, M, D, L, Lq, P, _lib.MSDA_F32,
                                 _lib.MSDA_PAD_ZEROS, 0, None)
    assert rc == ERR_SCRATCH_LIMIT
    assert b"MSDA_STATIC_SCRATCH_BYTES" in lib.msda_error_string(rc)
    # the largest call that fits is accepted as far as the argument checks (null pointers here)
    rc = lib.msda_forward_static(None, None, None, None, None, None, 64, S, M, D, L, 4096, P, _lib.MSDA_F32,
                                 _lib.MSDA_PAD_ZEROS, 0, None)
    assert rc == -1                       # MSDA_ERR_NULL_POINTER: past the limit check
    torch.cuda.synchronize()
