"""Shared test plumbing: paths, the `gpu` marker, golden loaders.

`-m "not gpu"`: oracle vs the golden vectors frozen from the reference's own CPU code, host-side
logic, the C-ABI library (loads, exports every declared symbol), gloo world_size=2 exchange.
`-m gpu`: the parity tests proper; they call the CUDA kernels through the C ABI (ctypes) and
compare with the oracle / golden vectors. Nothing here reads /root/reference at run time.
"""
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "gsplat-triton_b200")
GOLD = os.path.join(ROOT, "tests", "golden")
for p in (PKG, os.path.join(ROOT, "oracle"), ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)
os.environ.setdefault("GSPLAT_BACKEND", "b200")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    # a fresh checkout has no built library (it is git-ignored): build it once if the toolchain is here
    lib_path = os.path.join(PKG, "lib", "libgsplat_b200.so")
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    if not os.path.exists(lib_path) and os.path.exists(nvcc):
        import importlib.util

        spec = importlib.util.spec_from_file_location("_gsb_build", os.path.join(PKG, "build.py"))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        mod.build()


def pytest_collection_modifyitems(config, items):
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


def load_golden(name):
    """tests/golden/<name>.npz -> dict of torch tensors / python scalars."""
    data = np.load(os.path.join(GOLD, name + ".npz"))
    out = {}
    for k in data.files:
        v = data[k]
        out[k] = v.item() if v.ndim == 0 else torch.from_numpy(v)
    return out


@pytest.fixture(scope="session")
def golden():
    cache = {}

    def get(name):
        if name not in cache:
            cache[name] = load_golden(name)
        return cache[name]

    return get


def garden_scene(step=1, seed=42, device="cpu"):
    """Same protocol as the reference's gsplat/_helper.py:9-55 (load_test_data) on the committed fixture."""
    data = np.load(os.path.join(GOLD, "test_garden.npz"))
    height, width = int(data["height"]), int(data["width"])
    viewmats = torch.from_numpy(data["viewmats"]).float()
    Ks = torch.from_numpy(data["Ks"]).float()
    means = torch.from_numpy(data["means3d"]).float()
    colors = torch.from_numpy(data["colors"] / 255.0).float()
    sel = ((means >= -2) & (means <= 2)).all(dim=-1)
    means, colors = means[sel][::step], colors[sel][::step]
    g = torch.Generator().manual_seed(seed)
    N = means.shape[0]
    scales = torch.rand((N, 3), generator=g) * 0.02
    quats = torch.nn.functional.normalize(torch.randn((N, 4), generator=g), dim=-1)
    opacities = torch.rand((N,), generator=g)
    t = (means, quats, scales, opacities, colors, viewmats, Ks)
    return tuple(x.to(device) for x in t) + (width, height)
