"""CPU: the C-ABI library loads and exports every symbol include/gsplat_b200.h declares; the
product path refuses CPU tensors / a missing library instead of falling back."""
import ctypes
import os
import subprocess
import sys

import pytest
import torch

from conftest import PKG, ROOT

from gsplat.b200_impl import _lib


def test_header_parses_and_library_exports_every_symbol():
    protos = _lib.parse_header()
    expected = {
        "gsb_last_error", "gsb_abi_version", "gsb_launch_count", "gsb_reset_launch_count", "gsb_projection_fwd",
        "gsb_projection_bwd", "gsb_projection_packed_count", "gsb_projection_packed_write",
        "gsb_projection_packed_bwd", "gsb_sh_fwd", "gsb_sh_bwd", "gsb_isect_num_blocks", "gsb_isect_count",
        "gsb_scan_i32_to_i64", "gsb_isect_emit", "gsb_sort_pairs_temp_bytes", "gsb_sort_pairs", "gsb_isect_offsets",
        "gsb_rasterize_pack", "gsb_rasterize_fwd", "gsb_rasterize_bwd", "gsb_a2a_pack", "gsb_a2a_unpack",
    }
    assert expected <= set(protos), expected - set(protos)
    assert os.path.exists(_lib.LIB_PATH), "run `python gsplat-triton_b200/build.py` (or __graft_entry__.build())"
    dll = ctypes.CDLL(_lib.LIB_PATH)
    for name in protos:
        assert hasattr(dll, name), f"{name} declared in the header but not exported"
    dll.gsb_abi_version.restype = ctypes.c_int
    assert dll.gsb_abi_version() >= 1


def test_pure_queries_work_without_a_gpu():
    assert int(_lib.lib.gsb_isect_num_blocks(1000)) == 4
    assert int(_lib.lib.gsb_isect_num_blocks(0)) == 0
    _lib.lib.gsb_reset_launch_count()
    assert int(_lib.lib.gsb_launch_count()) == 0


def test_kernel_variant_switches_are_consistent():
    """GSB_OPT_* indices in the header are 0..GSB_OPT_COUNT-1, the library holds a default for each (the tuned path),
    set / get round-trips, out-of-range indices are refused, and every environment override `_lib.py` installs names an
    existing switch."""
    import re

    header = open(os.path.join(ROOT, "include", "gsplat_b200.h")).read()
    opts = {m.group(1): int(m.group(2)) for m in re.finditer(r"#define (GSB_OPT_[A-Z0-9_]+) (\d+)", header)}
    count = opts.pop("GSB_OPT_COUNT")
    assert sorted(opts.values()) == list(range(count)), opts
    lib = _lib.lib
    defaults = [int(lib.gsb_get_option(i)) for i in range(count)]
    assert all(d >= 0 for d in defaults), defaults
    assert defaults[opts["GSB_OPT_FWD_VARIANT"]] == 2 and defaults[opts["GSB_OPT_BWD_VARIANT"]] == 2
    assert defaults[opts["GSB_OPT_SH_STAGING"]] == 2 and defaults[opts["GSB_OPT_SH_FWD_STAGING"]] == 2  # cp.async
    assert defaults[opts["GSB_OPT_FINALIZE_VEC"]] == 0 and defaults[opts["GSB_OPT_CHAINED_LAUNCH"]] == 37
    assert int(lib.gsb_get_option(count)) == -1 and int(lib.gsb_set_option(count, 1)) != 0
    which = opts["GSB_OPT_ISECT_BIN_RADIX"]
    try:
        assert int(lib.gsb_set_option(which, 1)) == 0 and int(lib.gsb_get_option(which)) == 1
    finally:
        lib.gsb_set_option(which, defaults[which])
    source = open(os.path.join(PKG, "gsplat", "b200_impl", "_lib.py")).read()
    env = re.findall(r"\((\d+), \"(GSPLAT_B200_[A-Z_]+)\"\)", source)
    assert len(env) >= 6 and all(int(i) < count for i, _ in env), env


def test_library_is_sm100a_only():
    out = subprocess.run(["cuobjdump", "--list-elf", _lib.LIB_PATH], capture_output=True, text=True)
    if out.returncode != 0:
        pytest.skip("cuobjdump unavailable")
    archs = {line.split(".")[-2] for line in out.stdout.splitlines() if ".cubin" in line}
    assert archs == {"sm_100a"}, archs


def test_cpu_tensors_are_refused_not_emulated():
    from gsplat.b200_impl import fully_fused_projection, isect_tiles, spherical_harmonics

    with pytest.raises(RuntimeError, match="no CPU path"):
        fully_fused_projection(torch.zeros(4, 3), None, torch.ones(4, 4), torch.ones(4, 3), torch.eye(4)[None],
                               torch.eye(3)[None], 32, 32)
    with pytest.raises(RuntimeError, match="no CPU path"):
        spherical_harmonics(1, torch.zeros(4, 3), torch.zeros(4, 4, 3))
    with pytest.raises(RuntimeError, match="no CPU path"):
        isect_tiles(torch.zeros(1, 4, 2), torch.ones(1, 4, dtype=torch.int32), torch.ones(1, 4), 16, 2, 2)


def test_missing_library_fails_loudly():
    code = (
        "import sys; sys.path.insert(0, %r)\n"
        "from gsplat.b200_impl import _lib\n"
        "try:\n    _lib.lib.gsb_abi_version()\nexcept RuntimeError as e:\n    print('RAISED', 'no CPU fallback' in str(e))\n"
    ) % PKG
    env = dict(os.environ, GSPLAT_B200_LIB="/nonexistent/libgsplat_b200.so")
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=env)
    assert "RAISED True" in out.stdout, out.stdout + out.stderr


def test_unknown_backend_raises_like_the_reference():
    code = "import sys; sys.path.insert(0, %r)\nimport gsplat.rendering\n" % PKG
    env = dict(os.environ, GSPLAT_BACKEND="nope")
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=env)
    assert out.returncode != 0 and "Unknown GSPLAT_BACKEND" in out.stderr


def test_product_never_imports_the_oracle():
    for dirpath, _, files in os.walk(PKG):
        for f in files:
            if f.endswith(".py"):
                text = open(os.path.join(dirpath, f)).read()
                assert "gs_oracle" not in text and "import oracle" not in text, os.path.join(dirpath, f)
    assert os.path.isdir(os.path.join(ROOT, "oracle"))


def test_optimizer_ops_refuse_cpu_tensors():
    """SelectiveAdam / compute_relocation have no CPU path either: CPU tensors raise instead of falling back."""
    import pytest
    import torch
    from gsplat.optimizers import SelectiveAdam
    from gsplat.relocation import compute_relocation

    p = torch.zeros(4, 3, requires_grad=True)
    p.grad = torch.ones(4, 3)
    opt = SelectiveAdam([{"params": [p], "lr": 1e-3}], eps=1e-15, betas=(0.9, 0.999))
    with pytest.raises(RuntimeError, match="CUDA"):
        opt.step(visibility=torch.ones(4, dtype=torch.bool))
    with pytest.raises(RuntimeError, match="CUDA"):
        compute_relocation(torch.rand(4), torch.rand(4, 3), torch.ones(4, dtype=torch.int64), torch.ones(5, 5))


def test_rasterization_argument_validation_matches_the_reference():
    """Shape / mode checks fire before anything touches the GPU, with the reference's conditions and messages
    (rendering.py:242-293); valid CPU arguments then hit the no-CPU-path error, not a fallback."""
    import pytest
    import torch
    from gsplat.rendering import rasterization

    N, C = 8, 2
    ok = dict(means=torch.zeros(N, 3), quats=torch.ones(N, 4), scales=torch.ones(N, 3), opacities=torch.ones(N),
              colors=torch.ones(N, 3), viewmats=torch.eye(4).expand(C, 4, 4), Ks=torch.eye(3).expand(C, 3, 3),
              width=32, height=32)

    def call(**over):
        a = dict(ok, **over)
        return rasterization(a.pop("means"), a.pop("quats"), a.pop("scales"), a.pop("opacities"), a.pop("colors"),
                             a.pop("viewmats"), a.pop("Ks"), a.pop("width"), a.pop("height"), **a)

    for bad in (dict(means=torch.zeros(N, 2)), dict(quats=torch.ones(N, 3)), dict(scales=torch.ones(N + 1, 3)),
                dict(opacities=torch.ones(N, 1)), dict(viewmats=torch.eye(4).expand(C, 4, 4)[:, :3]),
                dict(Ks=torch.eye(3).expand(C + 1, 3, 3)), dict(render_mode="RGBD"), dict(colors=torch.ones(N + 1, 3)),
                dict(colors=torch.ones(N, 4, 3), sh_degree=2), dict(covars=torch.ones(N, 3, 2))):
        with pytest.raises(AssertionError):
            call(**bad)
    with pytest.raises(AssertionError, match="AbsGrad is not supported in distributed mode"):
        call(absgrad=True, distributed=True)
    with pytest.raises(AssertionError, match="per-Gaussian colors"):
        call(colors=torch.ones(C, N, 3), distributed=True)
    with pytest.raises(RuntimeError, match="CUDA"):
        call()


def test_stage_functions_validate_like_the_reference_wrappers():
    """The five backend functions check their arguments where the reference wrappers do (triton_impl/_wrapper.py:
    225-277, 511-523; cuda/_wrapper.py:260-300, 505-530) -- before any GPU work, so it can be seen on the CPU."""
    import pytest
    import torch
    from gsplat.b200_impl import fully_fused_projection, isect_tiles, rasterize_to_pixels, spherical_harmonics

    N, C = 6, 2
    means, quats, scales = torch.zeros(N, 3), torch.ones(N, 4), torch.ones(N, 3)
    viewmats, Ks = torch.eye(4).expand(C, 4, 4), torch.eye(3).expand(C, 3, 3)
    with pytest.raises(AssertionError):
        fully_fused_projection(means[:, :2], None, quats, scales, viewmats, Ks, 32, 32)
    with pytest.raises(AssertionError, match="covars or quats is required"):
        fully_fused_projection(means, None, None, scales, viewmats, Ks, 32, 32)
    with pytest.raises(AssertionError):
        fully_fused_projection(means, torch.ones(N, 3, 3), None, None, viewmats, Ks, 32, 32)  # covars are [N, 6] here
    with pytest.raises(AssertionError, match="sparse_grad is only supported when packed is True"):
        fully_fused_projection(means, None, quats, scales, viewmats, Ks, 32, 32, sparse_grad=True)
    with pytest.raises((NotImplementedError, ValueError, AssertionError)):
        fully_fused_projection(means, None, quats, scales, viewmats, Ks, 32, 32, camera_model="equirectangular")
    means2d, conics, colors, opac = torch.zeros(C, N, 2), torch.ones(C, N, 3), torch.ones(C, N, 3), torch.ones(C, N)
    offsets, flat = torch.zeros(C, 2, 2, dtype=torch.int32), torch.zeros(0, dtype=torch.int32)
    with pytest.raises(AssertionError):
        rasterize_to_pixels(means2d, conics[:, :, :2], colors, opac, 32, 32, 16, offsets, flat)
    with pytest.raises(AssertionError):
        rasterize_to_pixels(means2d, conics, colors, opac, 32, 32, 16, offsets, flat, backgrounds=torch.ones(C, 4))
    with pytest.raises(AssertionError, match="Assert Failed"):
        rasterize_to_pixels(means2d, conics, colors, opac, 64, 32, 16, offsets, flat)  # tile grid smaller than image
    with pytest.raises(ValueError, match="Unsupported number of color channels"):
        rasterize_to_pixels(means2d, conics, torch.ones(C, N, 600), opac, 32, 32, 16, offsets, flat)
    with pytest.raises(AssertionError):
        isect_tiles(means2d, torch.ones(C, N + 1, dtype=torch.int32), torch.ones(C, N), 16, 2, 2)
    with pytest.raises(AssertionError):
        spherical_harmonics(3, torch.ones(N, 3), torch.ones(N, 9, 3))  # (degree + 1)^2 coefficients needed
