"""TEST INFRASTRUCTURE -- loader for the reference's own CUDA op built by
oracle/build_ref.py into oracle/_ref/ (pybind module exporting
ms_deform_attn_forward / ms_deform_attn_backward, models/ops/src/vision.cpp:13-16)."""
import importlib.machinery
import importlib.util
import os

NAME = "MultiScaleDeformableAttention_ref"
PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", NAME + ".so")
_mod = None


def available():
    return os.path.exists(PATH)


def load():
    global _mod
    if _mod is None:
        import torch  # noqa: F401  (libtorch must be loaded first)
        loader = importlib.machinery.ExtensionFileLoader(NAME, PATH)
        spec = importlib.util.spec_from_loader(NAME, loader)
        _mod = importlib.util.module_from_spec(spec)
        loader.exec_module(_mod)
    return _mod
