"""TEST INFRASTRUCTURE -- numpy/ctypes front-end of oracle/msda_oracle.c.

forward(value, shapes, lsi, loc, aw, padding_mode)  -> out [N, Lq, M*D]
backward(..., grad_out, padding_mode)               -> (g_value, g_loc, g_aw)
Accepts numpy arrays or torch CPU tensors (float32 or float64); the arithmetic
type is the input dtype (bf16 tests round their inputs and pass float32).
See msda_oracle.c for the reference file:line each step restates.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libmsda_oracle.so")
_lib = None


def build():
    subprocess.run(["make", "-C", _HERE, "libmsda_oracle.so"], check=True,
                   stdout=subprocess.DEVNULL)
    return LIB_PATH


def _load():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        lib = ctypes.CDLL(LIB_PATH)
        vp, lg = ctypes.c_void_p, ctypes.c_long
        for sfx in ("f32", "f64"):
            f = getattr(lib, "msda_oracle_forward_" + sfx)
            f.restype = None
            f.argtypes = [vp] * 6 + [lg] * 7 + [ctypes.c_int]
            b = getattr(lib, "msda_oracle_backward_" + sfx)
            b.restype = None
            b.argtypes = [vp] * 9 + [lg] * 7 + [ctypes.c_int]
        _lib = lib
    return _lib


def _np(x, dtype=None):
    if hasattr(x, "detach"):
        x = x.detach().cpu()
        if str(x.dtype) == "torch.bfloat16":
            x = x.float()
        x = x.numpy()
    x = np.ascontiguousarray(x)
    return x if dtype is None else np.ascontiguousarray(x.astype(dtype, copy=False))


def _prep(value, shapes, lsi, loc, aw):
    value = _np(value)
    if value.dtype not in (np.float32, np.float64):
        raise TypeError("oracle works in float32 or float64")
    dt = value.dtype
    loc, aw = _np(loc, dt), _np(aw, dt)
    shapes, lsi = _np(shapes, np.int64), _np(lsi, np.int64)
    N, S, M, D = value.shape
    L = shapes.shape[0]
    Lq, P = loc.shape[1], loc.shape[4]
    assert shapes.shape == (L, 2) and lsi.shape == (L,)
    assert loc.shape == (N, Lq, M, L, P, 2) and aw.shape == (N, Lq, M, L, P)
    return value, shapes, lsi, loc, aw, (N, S, M, D, L, Lq, P), ("f32" if dt == np.float32 else "f64")


def _ptr(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def forward(value, shapes, lsi, loc, aw, padding_mode="zeros"):
    value, shapes, lsi, loc, aw, dims, sfx = _prep(value, shapes, lsi, loc, aw)
    N, S, M, D, L, Lq, P = dims
    out = np.empty((N, Lq, M * D), dtype=value.dtype)
    getattr(_load(), "msda_oracle_forward_" + sfx)(
        _ptr(value), _ptr(shapes), _ptr(lsi), _ptr(loc), _ptr(aw), _ptr(out), *dims,
        int(padding_mode == "border"))
    return out


def backward(value, shapes, lsi, loc, aw, grad_out, padding_mode="zeros"):
    value, shapes, lsi, loc, aw, dims, sfx = _prep(value, shapes, lsi, loc, aw)
    N, S, M, D, L, Lq, P = dims
    grad_out = _np(grad_out, value.dtype).reshape(N, Lq, M * D)
    gv, gl, ga = np.empty_like(value), np.empty_like(loc), np.empty_like(aw)
    getattr(_load(), "msda_oracle_backward_" + sfx)(
        _ptr(value), _ptr(shapes), _ptr(lsi), _ptr(loc), _ptr(aw), _ptr(grad_out),
        _ptr(gv), _ptr(gl), _ptr(ga), *dims, int(padding_mode == "border"))
    return gv, gl, ga
