"""ctypes binding + autograd wrapper of the C compositor oracle (oracle/gs_compositor.c).

TEST INFRASTRUCTURE, NOT A PRODUCT PATH (see the header of gs_compositor.c). `build()` compiles the C file with
gcc -O2 -fopenmp into oracle/_build/libgs_oracle.so (git-ignored, travels to the GPU box); `__graft_entry__.build()`
calls it, and it is also built on first use when gcc is present.
"""
import ctypes
import os
import subprocess
from typing import Optional

import torch
from torch import Tensor

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "gs_compositor.c")
LIB = os.path.join(HERE, "_build", "libgs_oracle.so")
_dll = None


def build(force: bool = False) -> str:
    if force or not os.path.exists(LIB) or os.path.getmtime(LIB) < os.path.getmtime(SRC):
        os.makedirs(os.path.dirname(LIB), exist_ok=True)
        # no -ffast-math: IEEE float32 arithmetic and glibc expf
        # the distribution gcc carries libgomp; the image's $CC (/opt/gcc) does not
        cc = os.environ.get("GSO_CC") or ("/usr/bin/gcc" if os.path.exists("/usr/bin/gcc") else "gcc")
        cmd = [cc, "-O2", "-fopenmp", "-shared", "-fPIC", "-std=c11", SRC, "-o", LIB, "-lm"]
        subprocess.run(cmd, check=True)
    return LIB


def _lib():
    global _dll
    if _dll is None:
        _dll = ctypes.CDLL(build())
        _dll.gso_rasterize_fwd.restype = ctypes.c_int
        _dll.gso_rasterize_bwd.restype = ctypes.c_int
    return _dll


def _p(t: Optional[Tensor]):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def _f(t: Optional[Tensor]):
    return None if t is None else t.detach().to("cpu", torch.float32).contiguous()


def rasterize_fwd(means2d, conics, colors, opacities, W, H, tile_size, isect_offsets, flatten_ids, backgrounds=None,
                  masks=None):
    """-> (render_colors [C,H,W,D], render_alphas [C,H,W,1], last_ids [C,H,W] int32, margin [C,H,W])."""
    C, th, tw = isect_offsets.shape
    D = colors.shape[-1]
    m, cn, col, op, bg = _f(means2d), _f(conics), _f(colors), _f(opacities), _f(backgrounds)
    offs = isect_offsets.detach().cpu().int().contiguous()
    fid = flatten_ids.detach().cpu().int().contiguous()
    m8 = masks.detach().cpu().to(torch.uint8).contiguous() if masks is not None else None
    rc = torch.empty((C, H, W, D), dtype=torch.float32)
    ra = torch.empty((C, H, W, 1), dtype=torch.float32)
    last = torch.empty((C, H, W), dtype=torch.int32)
    margin = torch.empty((C, H, W), dtype=torch.float32)
    rcode = _lib().gso_rasterize_fwd(
        ctypes.c_int(C), ctypes.c_int64(fid.numel()), ctypes.c_int(D), _p(m), _p(cn), _p(col), _p(op), _p(bg), _p(m8),
        ctypes.c_int(W), ctypes.c_int(H), ctypes.c_int(tile_size), ctypes.c_int(tw), ctypes.c_int(th), _p(offs),
        _p(fid), _p(rc), _p(ra), _p(last), _p(margin))
    assert rcode == 0
    return rc, ra, last, margin


def rasterize_bwd(means2d, conics, colors, opacities, W, H, tile_size, isect_offsets, flatten_ids, render_alphas,
                  last_ids, v_render_colors, v_render_alphas, backgrounds=None, masks=None, pixel_weight=None,
                  absgrad=False):
    """-> dict(v_means2d, v_conics, v_colors, v_opacities, v_backgrounds|None, v_means2d_abs|None), shaped like the
    primals. `pixel_weight` (bool [C,H,W]): False leaves the pixel out of the loss."""
    C, th, tw = isect_offsets.shape
    D = colors.shape[-1]
    m, cn, col, op, bg = _f(means2d), _f(conics), _f(colors), _f(opacities), _f(backgrounds)
    P = op.numel()
    offs = isect_offsets.detach().cpu().int().contiguous()
    fid = flatten_ids.detach().cpu().int().contiguous()
    m8 = masks.detach().cpu().to(torch.uint8).contiguous() if masks is not None else None
    pw = pixel_weight.detach().cpu().to(torch.uint8).contiguous() if pixel_weight is not None else None
    ra, last = _f(render_alphas), last_ids.detach().cpu().int().contiguous()
    vrc, vra = _f(v_render_colors), _f(v_render_alphas)
    out = dict(v_means2d=torch.empty_like(m), v_conics=torch.empty_like(cn), v_colors=torch.empty_like(col),
               v_opacities=torch.empty_like(op),
               v_backgrounds=torch.empty_like(bg) if bg is not None else None,
               v_means2d_abs=torch.empty_like(m) if absgrad else None)
    rcode = _lib().gso_rasterize_bwd(
        ctypes.c_int(C), ctypes.c_int64(P), ctypes.c_int64(fid.numel()), ctypes.c_int(D), _p(m), _p(cn), _p(col),
        _p(op), _p(bg), _p(m8), ctypes.c_int(W), ctypes.c_int(H), ctypes.c_int(tile_size), ctypes.c_int(tw),
        ctypes.c_int(th), _p(offs), _p(fid), _p(ra), _p(last), _p(vrc), _p(vra), _p(pw), _p(out["v_means2d"]),
        _p(out["v_conics"]), _p(out["v_colors"]), _p(out["v_opacities"]), _p(out["v_means2d_abs"]),
        _p(out["v_backgrounds"]))
    assert rcode == 0
    return out


class _Composite(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means2d, conics, colors, opacities, backgrounds, masks, W, H, tile_size, isect_offsets,
                flatten_ids, pixel_weight):
        rc, ra, last, margin = rasterize_fwd(means2d, conics, colors, opacities, W, H, tile_size, isect_offsets,
                                             flatten_ids, backgrounds, masks)
        ctx.save_for_backward(means2d, conics, colors, opacities, backgrounds, masks, isect_offsets, flatten_ids, ra,
                              last, pixel_weight)
        ctx.geom = (W, H, tile_size)
        ctx.mark_non_differentiable(last, margin)
        return rc, ra, last, margin

    @staticmethod
    def backward(ctx, v_rc, v_ra, _v_last, _v_margin):
        (means2d, conics, colors, opacities, backgrounds, masks, isect_offsets, flatten_ids, ra, last,
         pixel_weight) = ctx.saved_tensors
        W, H, tile_size = ctx.geom
        g = rasterize_bwd(means2d, conics, colors, opacities, W, H, tile_size, isect_offsets, flatten_ids, ra, last,
                          v_rc, v_ra, backgrounds, masks, pixel_weight)
        return (g["v_means2d"].reshape(means2d.shape), g["v_conics"].reshape(conics.shape),
                g["v_colors"].reshape(colors.shape), g["v_opacities"].reshape(opacities.shape),
                g["v_backgrounds"]) + (None,) * 7


def rasterize_to_pixels(means2d, conics, colors, opacities, image_width, image_height, tile_size, isect_offsets,
                        flatten_ids, backgrounds=None, masks=None, return_last_ids=False, return_margin=False,
                        pixel_weight=None):
    """Same contract as gs_oracle.rasterize_to_pixels (differentiable, CPU float32 tensors), on the C compositor.
    `return_margin`: also return the per-pixel decision margin (see gs_compositor.c)."""
    rc, ra, last, margin = _Composite.apply(means2d, conics, colors, opacities, backgrounds, masks, image_width,
                                            image_height, tile_size, isect_offsets, flatten_ids, pixel_weight)
    out = (rc, ra)
    if return_last_ids:
        out += (last,)
    if return_margin:
        out += (margin,)
    return out
