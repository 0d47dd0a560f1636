"""Guard bands around every output of the C ABI (compute-sanitizer is closed on this pool).

Each output (out, grad_value, grad_sampling_loc, grad_attn_weight, workspace) is a slice in
the middle of a larger allocation filled with a byte pattern; after the launches the bands on
both sides must be bitwise untouched, and the outputs themselves must hold no byte of the
pattern where a value is expected (every element written).  Runs every kernel family on
shapes that exercise ragged ends: Lq not a multiple of the 16-query unit, fewer than 4
levels, a single query, blocks that fall back inside a window launch.
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

import mfl_b200  # noqa: E402,F401
from mfl_b200 import _lib, synthetic  # noqa: E402

DEV = "cuda"
GUARD = 4096          # bytes each side
PATTERN = 0xA5


class Guarded:
    """A tensor of `shape` carved out of the middle of a byte buffer full of PATTERN."""

    def __init__(self, shape, dtype):
        self.nbytes = int(np.prod(shape)) * torch.empty((), dtype=dtype).element_size()
        self.raw = torch.full((GUARD + self.nbytes + GUARD,), PATTERN, dtype=torch.uint8, device=DEV)
        self.t = self.raw[GUARD:GUARD + self.nbytes].view(dtype).view(shape)

    def bands_intact(self):
        return bool((self.raw[:GUARD] == PATTERN).all()) and bool((self.raw[GUARD + self.nbytes:] == PATTERN).all())

    def fully_written(self):
        """No element still carries the fill pattern (0xA5A5.. is a finite but absurd number)."""
        return not bool((self.raw[GUARD:GUARD + self.nbytes].view(-1, self.t.element_size()) == PATTERN).all(-1).any())


CASES = [
    # N, T, M, D, Lq (None = encoder tokens), dist
    (2, [48, 24, 12, 6], 2, 64, None, "model"),
    (1, [48, 24, 12, 6], 2, 64, None, "uniform"),
    (2, [30, 15], 3, 32, None, "model"),
    (1, [20, 10, 5, 3], 2, 64, 1, "model"),
    (1, [40, 20, 10, 5], 1, 64, 21, "spill"),
    (1, [33], 2, 16, 19, "spill"),
]
FLAGS = [0, _lib.MSDA_FLAG_FORCE_GENERIC, _lib.MSDA_FLAG_NO_WINDOW, _lib.MSDA_FLAG_FORCE_WINDOW]


@pytest.mark.parametrize("case", CASES, ids=[f"c{i}" for i in range(len(CASES))])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16], ids=["f32", "bf16"])
def test_outputs_stay_inside_their_buffers(case, dtype):
    N, T, M, D, Lq, dist = case
    enc = Lq is None
    inp = synthetic.make_inputs(N, sum(T) if enc else Lq, M, D, 4, [(1, t) for t in T], dtype, DEV, seed=21,
                                dist=dist, value="normal", enc_ref=enc)
    lib = _lib.load()
    v, loc, aw = inp["value"], inp["loc"], inp["aw"]
    S, L, Lq_, P = v.shape[1], inp["shapes"].shape[0], loc.shape[1], loc.shape[4]
    dims = (N, S, M, D, L, Lq_, P)
    dt = {torch.float32: 0, torch.bfloat16: 2}[dtype]
    st = torch.cuda.current_stream().cuda_stream
    ws_bytes = lib.msda_backward_workspace_bytes(*dims, dt)
    for padding in (0, 1):
        for fl in FLAGS:
            out = Guarded((N, Lq_, M * D), dtype)
            gv, gl, ga = Guarded(v.shape, dtype), Guarded(loc.shape, dtype), Guarded(aw.shape, dtype)
            ws = Guarded((max(ws_bytes // 4, 1),), torch.float32)
            _lib.check(lib.msda_forward(v.data_ptr(), inp["shapes"].data_ptr(), inp["lsi"].data_ptr(),
                                        loc.data_ptr(), aw.data_ptr(), out.t.data_ptr(), *dims, dt, padding, fl, st))
            _lib.check(lib.msda_backward(v.data_ptr(), inp["shapes"].data_ptr(), inp["lsi"].data_ptr(),
                                         loc.data_ptr(), aw.data_ptr(), inp["grad_out"].data_ptr(),
                                         gv.t.data_ptr(), gl.t.data_ptr(), ga.t.data_ptr(), *dims, dt, padding, fl,
                                         ws.t.data_ptr() if ws_bytes else None, ws_bytes, st))
            torch.cuda.synchronize()
            for name, g in (("out", out), ("grad_value", gv), ("grad_loc", gl), ("grad_aw", ga), ("workspace", ws)):
                assert g.bands_intact(), (name, padding, fl)
            for name, g in (("out", out), ("grad_value", gv), ("grad_loc", gl), ("grad_aw", ga)):
                assert g.fully_written(), (name, padding, fl)
                assert bool(torch.isfinite(g.t.float()).all()), (name, padding, fl)


def test_bf16_backward_without_queries_zeroes_grad_value():
    """Lq == 0 with bf16: grad_value is not the accumulation image, it must still come back as zeros
    (ADVICE round 1)."""
    import torch
    from mfl_b200 import MultiScaleDeformableAttention as MSDA
    from mfl_b200 import synthetic
    inp = synthetic.make_inputs(2, 0, 2, 64, 4, [(1, 20), (1, 10)], torch.bfloat16, "cuda", seed=3)
    gv = torch.full_like(inp["value"], 7.0)
    out = MSDA.ms_deform_attn_backward(inp["value"], inp["shapes"], inp["lsi"], inp["loc"], inp["aw"],
                                       inp["grad_out"], 64, grad_value=gv)
    torch.cuda.synchronize()
    assert float(out[0].abs().max()) == 0.0 and out[1].numel() == 0 and out[2].numel() == 0
