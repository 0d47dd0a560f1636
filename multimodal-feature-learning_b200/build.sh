#!/bin/bash
# Build libmsda_b200.so (sm_100a only) next to this script.  nvcc cross-compiles
# without a GPU.  Usage: ./build.sh [extra nvcc flags]
set -euo pipefail
here="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 \
     -shared -Xcompiler -fPIC "$@" \
     -o "$here/libmsda_b200.so" "$here/csrc/msda_capi.cu"
