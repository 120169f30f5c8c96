"""ctypes binding of libgsplat_b200.so (the C ABI declared in include/gsplat_b200.h).

The prototypes are read from the header itself, so the Python side cannot drift from the ABI.
There is NO fallback: if the library is missing or a call fails, a RuntimeError is raised.
"""
import ctypes
import os
import re
from typing import Optional

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_PKG_ROOT = os.path.dirname(os.path.dirname(_HERE))  # gsplat-triton_b200/
_REPO_ROOT = os.path.dirname(_PKG_ROOT)
LIB_PATH = os.environ.get("GSPLAT_B200_LIB", os.path.join(_PKG_ROOT, "lib", "libgsplat_b200.so"))
HEADER_PATH = os.path.join(_REPO_ROOT, "include", "gsplat_b200.h")

_CTYPES = {
    "int": ctypes.c_int,
    "int64_t": ctypes.c_int64,
    "float": ctypes.c_float,
    "void": None,
    "cudaStream_t": ctypes.c_void_p,
}


def parse_header(path: str = HEADER_PATH):
    """Returns {name: (restype, [(argtype, argname), ...])} for every prototype in the header."""
    text = open(path).read()
    text = re.sub(r"/\*.*?\*/", " ", text, flags=re.S)
    text = re.sub(r"^\s*#.*$", " ", text, flags=re.M)
    protos = {}
    for m in re.finditer(r"([A-Za-z_][\w\s\*]*?)\b(gsb_\w+)\s*\(([^)]*)\)\s*;", text):
        ret, name, args = m.group(1).strip(), m.group(2), m.group(3).strip()
        if "*" in ret:
            restype = ctypes.c_char_p if "char" in ret else ctypes.c_void_p
        else:
            restype = _CTYPES[ret.replace("const", "").strip()]
        argl = []
        if args and args != "void":
            for a in args.split(","):
                a = a.strip()
                if "*" in a:
                    argl.append((ctypes.c_void_p, a.split("*")[-1].strip()))
                else:
                    toks = a.replace("const", "").split()
                    argl.append((_CTYPES[toks[0]], toks[-1]))
        protos[name] = (restype, argl)
    return protos


class _Timing:
    """Optional per-entry-point CUDA-event timing (bench.py's per-stage roofline): when enabled, every
    status-returning launcher is bracketed by events recorded on the stream it launches on."""

    def __init__(self):
        self.enabled = False
        self.records = {}

    def reset(self, enabled: bool):
        self.enabled = enabled
        self.records = {}

    def summary(self):
        """{entry point: (calls, total ms)}; synchronises."""
        torch.cuda.synchronize()
        return {k: (len(v), sum(a.elapsed_time(b) for a, b in v)) for k, v in self.records.items()}


timing = _Timing()


class _Lib:
    def __init__(self):
        self._dll = None
        self.protos = None

    def load(self):
        if self._dll is not None:
            return self._dll
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"gsplat b200 backend: {LIB_PATH} not found. Build it with "
                f"`python {os.path.join(_PKG_ROOT, 'build.py')}` (needs nvcc); there is no CPU fallback."
            )
        dll = ctypes.CDLL(LIB_PATH)
        self.protos = parse_header()
        for name, (restype, argl) in self.protos.items():
            fn = getattr(dll, name)  # AttributeError if the symbol is missing
            fn.restype = restype
            fn.argtypes = [t for t, _ in argl]
        self._dll = dll
        # kernel-variant switches for A/B runs (include/gsplat_b200.h: GSB_OPT_*); unset = the tuned defaults
        for which, env in ((0, "GSPLAT_B200_FWD_VARIANT"), (1, "GSPLAT_B200_BWD_VARIANT"),
                           (3, "GSPLAT_B200_SH_STAGING"), (4, "GSPLAT_B200_FINALIZE_VEC"),
                           (5, "GSPLAT_B200_SH_CTAS"), (6, "GSPLAT_B200_SH_FWD_STAGING"),
                           (7, "GSPLAT_B200_SH_FWD_CTAS"), (8, "GSPLAT_B200_CHAINED_LAUNCH")):
            if os.environ.get(env):
                dll.gsb_set_option(which, int(os.environ[env]))
        return dll

    def __getattr__(self, name):
        dll = self.load()
        fn = getattr(dll, name)
        restype, argl = self.protos[name]
        if restype is not ctypes.c_int or not argl or argl[-1][1] != "stream":
            return fn  # plain query, not a status-returning launcher

        def checked(*args):
            if timing.enabled:
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                rc = fn(*args)
                b.record()
                timing.records.setdefault(name, []).append((a, b))
            else:
                rc = fn(*args)
            if rc != 0:
                raise RuntimeError(f"{name} failed ({rc}): {dll.gsb_last_error().decode()}")
            return rc

        checked.__name__ = name
        setattr(self, name, checked)
        return checked


lib = _Lib()


def ptr(t: Optional[torch.Tensor]):
    """Device pointer of a contiguous tensor (None -> NULL)."""
    if t is None:
        return None
    assert t.is_contiguous(), "b200 backend expects contiguous tensors"
    return t.data_ptr()


_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)


def stream_of(t: torch.Tensor):
    """cudaStream_t of the current stream of the tensor's device (the raw getter: the `torch.cuda.current_stream`
    object costs ~10 us of Python per call, a dozen calls per step)."""
    if _raw_stream is None:
        return torch.cuda.current_stream(t.device).cuda_stream
    idx = t.device.index
    return _raw_stream(torch.cuda.current_device() if idx is None else idx)


def require_cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise RuntimeError(
                "gsplat b200 backend: tensors must live on a CUDA device (sm_100a); there is no CPU path"
            )
