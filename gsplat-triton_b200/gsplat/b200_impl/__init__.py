"""B200 (sm_100a) backend of gsplat's rasterization() hot path.

Drop-in sibling of ``gsplat/triton_impl`` and ``gsplat/cuda`` in the reference: the five
functions of ``_wrapper.py`` are what ``gsplat/rendering.py`` imports under ``GSPLAT_BACKEND=b200``.
"""
from ._accumulate import accumulate, rasterize_to_pixels_torch  # noqa: F401
from ._wrapper import (  # noqa: F401
    fully_fused_projection,
    isect_offset_encode,
    isect_tiles,
    rasterize_to_indices_in_range,
    rasterize_to_pixels,
    spherical_harmonics,
)
