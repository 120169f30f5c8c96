"""gsplat (b200 build): `rasterization()` and its five backend stages on hand-written sm_100a kernels.

Only the hot path of the reference is provided: `gsplat.rendering.rasterization`,
`gsplat.distributed`, `gsplat.b200_impl`, plus `gsplat.optimizers.SelectiveAdam` (section 8f). For parity runs against the reference's own backends set
GSPLAT_REFERENCE_PATH=<checkout of hieu1999210/gsplat-triton>: its `gsplat/` directory is appended to this
package's search path so `gsplat.triton_impl`, `gsplat.cuda`, `gsplat._helper` resolve there while
`gsplat.rendering`, `gsplat.distributed`, `gsplat.b200_impl` stay ours.
"""
import os as _os

_ref = _os.environ.get("GSPLAT_REFERENCE_PATH")
if _ref:
    _ref_pkg = _os.path.join(_ref, "gsplat")
    if _os.path.isdir(_ref_pkg) and _ref_pkg not in __path__:
        __path__.append(_ref_pkg)

from .version import __version__  # noqa: E402,F401


def __getattr__(name):
    # lazy: importing gsplat must not require a GPU or the built library
    if name in ("rasterization",):
        from .rendering import rasterization

        return rasterization
    if name in ("fully_fused_projection", "isect_offset_encode", "isect_tiles", "rasterize_to_pixels",
                "spherical_harmonics"):
        from . import rendering

        return getattr(rendering, name)
    if name == "rasterize_to_indices_in_range":
        from .b200_impl import rasterize_to_indices_in_range

        return rasterize_to_indices_in_range
    if name == "accumulate":
        from .b200_impl import accumulate

        return accumulate
    if name == "quat_scale_to_covar_preci":
        from .b200_impl._covar import quat_scale_to_covar_preci

        return quat_scale_to_covar_preci
    if name in ("DefaultStrategy", "MCMCStrategy", "Strategy"):
        # the reference's own strategy package, unmodified, resolved through GSPLAT_REFERENCE_PATH (it is pure torch
        # on top of rasterization()'s `meta`, gsplat.relocation and gsplat.quat_scale_to_covar_preci)
        from . import strategy

        return getattr(strategy, name)
    if name == "SelectiveAdam":
        from .optimizers import SelectiveAdam

        return SelectiveAdam
    raise AttributeError(f"module 'gsplat' has no attribute {name!r}")
