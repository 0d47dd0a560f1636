"""TEST INFRASTRUCTURE -- build the REFERENCE's own CUDA op (unmodified kernels)
for sm_100a into oracle/_ref/ so the GPU tests and bench.py can use it as the
GPU-side oracle and as the kernel-to-beat.

Sources are compiled from /root/reference/models/ops/src where they lie
(vision.cpp, cpu/ms_deform_attn_cpu.cpp, cuda/ms_deform_attn_cuda.cu, which
includes cuda/ms_deform_im2col_cuda.cuh).  The only change is the one token
torch >= 2 requires -- `value.type()` -> `value.scalar_type()` inside the
AT_DISPATCH macro (cuda/ms_deform_attn_cuda.cu:64,134) -- applied with a text
substitution to a scratch copy under /tmp; no reference source enters the repo
and oracle/_ref/ is git-ignored.  The reference's setup.py is NOT used (it
refuses to build without a visible GPU, models/ops/setup.py:36-47); flags are the
ones it would pass (:40-45).

Load the result with oracle.ref_cuda.load() (needs a CUDA device to run).
"""
import os
import shutil
import sys
import tempfile

SRC = "/root/reference/models/ops/src"
HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")
NAME = "MultiScaleDeformableAttention_ref"


def main():
    if not os.path.isdir(SRC):
        sys.exit("reference sources not found at " + SRC)
    os.makedirs(OUT, exist_ok=True)
    if os.path.exists(os.path.join(OUT, NAME + ".so")) and "--force" not in sys.argv:
        print("oracle/_ref already built")
        return
    os.environ.setdefault("TORCH_CUDA_ARCH_LIST", "10.0")
    os.environ["CC"] = "/usr/bin/gcc" if os.path.exists("/usr/bin/gcc") else os.environ.get("CC", "gcc")
    os.environ["CXX"] = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else os.environ.get("CXX", "g++")
    from torch.utils import cpp_extension
    tmp = tempfile.mkdtemp(prefix="msda_ref_src_")
    try:
        work = os.path.join(tmp, "src")
        shutil.copytree(SRC, work)
        cu = os.path.join(work, "cuda", "ms_deform_attn_cuda.cu")
        text = open(cu).read()
        old = "AT_DISPATCH_FLOATING_TYPES(value.type(),"
        assert text.count(old) == 2
        open(cu, "w").write(text.replace(old, "AT_DISPATCH_FLOATING_TYPES(value.scalar_type(),"))
        build_dir = os.path.join(tmp, "build")
        os.makedirs(build_dir)
        cpp_extension.load(
            name=NAME,
            sources=[os.path.join(work, "vision.cpp"), os.path.join(work, "cpu", "ms_deform_attn_cpu.cpp"), cu],
            extra_include_paths=[work],
            extra_cflags=["-DWITH_CUDA", "-O2"],
            extra_cuda_cflags=["-DCUDA_HAS_FP16=1", "-D__CUDA_NO_HALF_OPERATORS__",
                               "-D__CUDA_NO_HALF_CONVERSIONS__", "-D__CUDA_NO_HALF2_OPERATORS__",
                               "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo"],
            with_cuda=True, build_directory=build_dir, is_python_module=False, verbose=False)
        shutil.copy(os.path.join(build_dir, NAME + ".so"), os.path.join(OUT, NAME + ".so"))
        print("built", os.path.join(OUT, NAME + ".so"))
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


if __name__ == "__main__":
    main()
