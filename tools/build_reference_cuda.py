"""Builds the UNMODIFIED reference CUDA backend (gsplat/cuda/csrc/*.cu|*.cpp + gsplat/cuda/ext.cpp) for sm_100 as
`baseline/_ref/gsplat/csrc.so`, the pre-built module the reference's `cuda/_backend.py:136` imports first.

Why a recipe of our own: as shipped, neither of the reference's two build routes works with this image's torch 2.11:
  * `setup.py` globs `gsplat/cuda/csrc/*` only, so `ext.cpp` (the PYBIND11_MODULE) is left out and the installed
    `gsplat/csrc.so` has no `PyInit_csrc` ("dynamic module does not define module export function");
  * the JIT route calls the private `torch.utils.cpp_extension._jit_compile` with the pre-2.6 argument list
    (`_backend.py:70`: TypeError, missing `with_sycl`).
Same sources, same flags as `_backend.py:150-156` (-O3 -use_fast_math), via the public `cpp_extension.load`.
Run in the build container (nvcc cross-compiles; no GPU needed): python tools/build_reference_cuda.py
"""
import glob
import os
import shutil
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.environ.get("GSPLAT_REF_SRC", "/root/reference")
OUT = os.path.join(ROOT, "baseline", "_ref", "gsplat")
BUILD = os.path.join(ROOT, "baseline", "_ref", "build_csrc")

os.environ.setdefault("TORCH_CUDA_ARCH_LIST", "10.0")
os.environ.setdefault("MAX_JOBS", "8")
from torch.utils.cpp_extension import load  # noqa: E402

cuda = os.path.join(SRC, "gsplat", "cuda")
sources = sorted(glob.glob(os.path.join(cuda, "csrc", "*.cu")) + glob.glob(os.path.join(cuda, "csrc", "*.cpp")))
sources.append(os.path.join(cuda, "ext.cpp"))
os.makedirs(BUILD, exist_ok=True)
load(
    name="csrc",
    sources=sources,
    extra_cflags=["-O3", "-Wno-attributes"],
    extra_cuda_cflags=["-O3", "-use_fast_math", "--expt-relaxed-constexpr", "-diag-suppress", "20012,186"],
    extra_include_paths=[os.path.join(cuda, "include"), os.path.join(cuda, "csrc", "third_party", "glm")],
    build_directory=BUILD,
    is_python_module=False,
    verbose="-v" in sys.argv,
)
shutil.copy(os.path.join(BUILD, "csrc.so"), os.path.join(OUT, "csrc.so"))
shutil.rmtree(BUILD)
print("built", os.path.join(OUT, "csrc.so"))
