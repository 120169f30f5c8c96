#!/bin/bash
# Installs the UNMODIFIED reference (hieu1999210/gsplat-triton, /root/reference) into baseline/_ref for the
# same-box comparison runs of tools/ref_backends.py. baseline/_ref is git-ignored but
# travels to the GPU box with gpurun. The reference's CUDA backend is cross-compiled here for sm_100 by the
# reference's own setup.py (TORCH_CUDA_ARCH_LIST=10.0); the Triton backend's one CUDA file (radix_sort.cu)
# is JIT-built by the reference itself at first import (torch.utils.cpp_extension.load).
set -euo pipefail
ROOT="$(cd "$(dirname "$0")/.." && pwd)"
SRC=/tmp/gsplat_ref_src
rm -rf "$SRC" && cp -r /root/reference "$SRC"
rm -rf "$ROOT/baseline/_ref" && mkdir -p "$ROOT/baseline/_ref"
cd "$SRC"
TORCH_CUDA_ARCH_LIST="10.0" MAX_JOBS="${MAX_JOBS:-8}" python -m pip install --no-index --no-build-isolation \
    --no-deps --find-links /opt/wheelhouse --target "$ROOT/baseline/_ref" "$SRC" 2>&1 | tail -20
# the stage tests and the bundled scene the reference's own tests load; the Triton backend's CUDA file, which the
# reference's MANIFEST leaves out of the wheel (its __init__ JIT-builds it at import)
cp -r "$SRC/triton_tests" "$SRC/tests" "$SRC/assets" "$SRC/examples" "$SRC/profiling" "$ROOT/baseline/_ref/"
cp "$SRC/gsplat/triton_impl/radix_sort/radix_sort.cu" "$ROOT/baseline/_ref/gsplat/triton_impl/radix_sort/"
# the csrc.so that setup.py produced lacks the pybind module (ext.cpp is not in its source glob): rebuild it
python "$ROOT/tools/build_reference_cuda.py"
# pre-build the Triton backend's radix_sort JIT module into a cache that travels with the snapshot
TORCH_EXTENSIONS_DIR="$ROOT/baseline/_ref/torch_ext" TORCH_CUDA_ARCH_LIST="10.0" PYTHONPATH="$ROOT/baseline/_ref" \
    python -c "from gsplat.triton_impl.radix_sort import radix_sort; print(radix_sort)"
ls "$ROOT/baseline/_ref" "$ROOT/baseline/_ref/gsplat" | head -40
