"""ctypes binding of libmsda_b200.so (the C ABI in include/msda_b200.h).

There is deliberately NO fallback: if the shared library is missing or a
non-CUDA tensor reaches the op, callers get an exception (the reference raises
"Not implemented on the CPU", models/ops/src/ms_deform_attn.h:38,60).
"""
import ctypes
import os
import subprocess
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libmsda_b200.so")
BUILD_SCRIPT = os.path.join(_HERE, "build.sh")

MSDA_F32, MSDA_F64, MSDA_BF16 = 0, 1, 2
MSDA_PAD_ZEROS, MSDA_PAD_BORDER = 0, 1
MSDA_FLAG_FORCE_GENERIC = 1
MSDA_FLAG_SKIP_ZERO_FILL = 2
MSDA_FLAG_NO_WINDOW = 4
MSDA_FLAG_FORCE_WINDOW = 8
MSDA_FLAG_DETERMINISTIC = 16
MSDA_FLAG_NO_TILE = 32
MSDA_FLAG_PLAN = 64
MSDA_FLAG_SAVE_PLAN = 128
MSDA_FLAG_PERSIST = 1 << 20   # (bits 8..19 are the QTILE override)
ABI_VERSION = 2


def flag_qtile(k):
    return (int(k) & 0xFFF) << 8

PADDING_MODES = {"zeros": MSDA_PAD_ZEROS, "border": MSDA_PAD_BORDER}

# every symbol include/msda_b200.h declares (checked by tests/test_abi.py)
EXPORTED_SYMBOLS = (
    "msda_abi_version",
    "msda_error_string",
    "msda_kernel_family",
    "msda_kernel_choice",
    "msda_forward",
    "msda_forward_workspace_bytes",
    "msda_forward_ex",
    "msda_forward_static",
    "msda_backward_workspace_bytes",
    "msda_backward_workspace_bytes_flags",
    "msda_backward",
    "msda_backward_ex",
    "msda_module_supported",
    "msda_module_forward",
    "msda_module_backward",
    "msda_frontend_forward",
    "msda_frontend_backward",
    "msda_dam_scatter",
    "msda_encoder_reference_points",
    "msda_prepare_encoder_inputs",
    "msda_token_gather",
    "msda_token_scatter",
    "msda_last_cuda_error",
    "msda_last_cuda_error_string",
)

_lock = threading.Lock()
_lib = None


def build(extra_flags=()):
    """Compile libmsda_b200.so for sm_100a with nvcc (works without a GPU)."""
    subprocess.run(["bash", BUILD_SCRIPT, *extra_flags], check=True)
    return LIB_PATH


def _declare(lib):
    c = ctypes
    i64, vp, u32 = c.c_int64, c.c_void_p, c.c_uint
    lib.msda_abi_version.restype = c.c_int
    lib.msda_abi_version.argtypes = []
    lib.msda_error_string.restype = c.c_char_p
    lib.msda_error_string.argtypes = [c.c_int]
    lib.msda_kernel_family.restype = c.c_char_p
    lib.msda_kernel_family.argtypes = [i64, c.c_int, u32]
    lib.msda_kernel_choice.restype = c.c_char_p
    lib.msda_kernel_choice.argtypes = [i64] * 7 + [c.c_int, u32]
    lib.msda_last_cuda_error.restype = c.c_int
    lib.msda_last_cuda_error.argtypes = []
    lib.msda_last_cuda_error_string.restype = c.c_char_p
    lib.msda_last_cuda_error_string.argtypes = []
    lib.msda_forward.restype = c.c_int
    lib.msda_forward.argtypes = [vp, vp, vp, vp, vp, vp] + [i64] * 7 + [c.c_int, c.c_int, u32, vp]
    lib.msda_forward_workspace_bytes.restype = c.c_size_t
    lib.msda_forward_workspace_bytes.argtypes = [i64] * 7 + [c.c_int, u32]
    lib.msda_forward_ex.restype = c.c_int
    lib.msda_forward_ex.argtypes = ([vp, vp, vp, vp, vp, vp] + [i64] * 7 +
                                    [c.c_int, c.c_int, u32, vp, c.c_size_t, vp])
    lib.msda_forward_static.restype = c.c_int
    lib.msda_forward_static.argtypes = [vp, vp, vp, vp, vp, vp] + [i64] * 7 + [c.c_int, c.c_int, u32, vp]
    lib.msda_backward_ex.restype = c.c_int
    lib.msda_backward_ex.argtypes = (
        [vp] * 9 + [i64] * 7 + [c.c_int, c.c_int, u32, vp, c.c_size_t, vp, vp]
    )
    lib.msda_backward_workspace_bytes.restype = c.c_size_t
    lib.msda_backward_workspace_bytes.argtypes = [i64] * 7 + [c.c_int]
    lib.msda_backward_workspace_bytes_flags.restype = c.c_size_t
    lib.msda_backward_workspace_bytes_flags.argtypes = [i64] * 7 + [c.c_int, u32]
    lib.msda_backward.restype = c.c_int
    lib.msda_backward.argtypes = (
        [vp] * 9 + [i64] * 7 + [c.c_int, c.c_int, u32, vp, c.c_size_t, vp]
    )
    lib.msda_module_supported.restype = c.c_int
    lib.msda_module_supported.argtypes = [i64] * 7 + [c.c_int, c.c_int, u32]
    lib.msda_module_forward.restype = c.c_int
    lib.msda_module_forward.argtypes = ([vp] * 8 + [i64] * 7 + [c.c_int, c.c_int, c.c_int, u32, vp, c.c_size_t, vp])
    lib.msda_module_backward.restype = c.c_int
    lib.msda_module_backward.argtypes = ([vp] * 12 + [i64] * 7 + [c.c_int, c.c_int, c.c_int, u32, vp, c.c_size_t,
                                                                 vp, vp])
    lib.msda_frontend_forward.restype = c.c_int
    lib.msda_frontend_forward.argtypes = [vp] * 6 + [i64] * 5 + [c.c_int, c.c_int, vp]
    lib.msda_frontend_backward.restype = c.c_int
    lib.msda_frontend_backward.argtypes = [vp] * 9 + [i64] * 5 + [c.c_int, c.c_int, vp]
    lib.msda_encoder_reference_points.restype = c.c_int
    lib.msda_encoder_reference_points.argtypes = [vp] * 3 + [i64] * 3 + [vp]
    lib.msda_prepare_encoder_inputs.restype = c.c_int
    lib.msda_prepare_encoder_inputs.argtypes = [vp] * 9 + [i64] * 3 + [c.c_int, vp]
    lib.msda_token_gather.restype = c.c_int
    lib.msda_token_gather.argtypes = [vp] * 4 + [i64] * 4 + [vp]
    lib.msda_token_scatter.restype = c.c_int
    lib.msda_token_scatter.argtypes = [vp] * 6 + [i64] * 4 + [c.c_int, vp]
    lib.msda_dam_scatter.restype = c.c_int
    lib.msda_dam_scatter.argtypes = [vp] * 5 + [i64] * 6 + [c.c_int, c.c_int, vp]
    return lib


def load():
    """Return the loaded library; raise if it has not been built."""
    global _lib
    if _lib is None:
        with _lock:
            if _lib is None:
                if not os.path.exists(LIB_PATH):
                    raise RuntimeError(
                        "libmsda_b200.so is not built (%s). Run `python -c 'import "
                        "__graft_entry__ as g; g.build()'` or %s. There is no CPU/PyTorch "
                        "fallback for MSDeformAttn in this package." % (LIB_PATH, BUILD_SCRIPT)
                    )
                _lib = _declare(ctypes.CDLL(LIB_PATH))
    return _lib


def check(status):
    """Turn a non-zero msda_status into a RuntimeError (c10::Error analogue)."""
    if status != 0:
        lib = load()
        msg = lib.msda_error_string(status).decode()
        if status == -7:
            msg += ": " + lib.msda_last_cuda_error_string().decode()
        raise RuntimeError("msda_b200: " + msg)
