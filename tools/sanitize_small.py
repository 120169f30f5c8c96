"""Small end-to-end pass for compute-sanitizer: packed and unpacked, SH, antialiased, absgrad, multi-camera,
speculative intersection buffers (second call), depth modes - every kernel of the path at a size memcheck can walk."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "gsplat-triton_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
os.environ.setdefault("GSPLAT_BACKEND", "b200")
from conftest import garden_scene  # noqa: E402

from gsplat.rendering import rasterization  # noqa: E402

dev = "cuda"
means, quats, scales, opac, colors, viewmats, Ks, W, H = garden_scene(step=200)
g = torch.Generator().manual_seed(0)
shs = torch.cat([colors[:, None, :], torch.randn(len(means), 8, 3, generator=g) * 0.2], dim=1)
n = 0
for packed in (False, True):
    for mode, model in (("classic", "pinhole"), ("antialiased", "fisheye")):
        for rep in range(2):  # the second call takes the speculative path
            p = [t.clone().to(dev).requires_grad_(True) for t in (means, quats, scales * 3 + 0.002, opac, shs)]
            rc, ra, meta = rasterization(*p, viewmats[:2].to(dev), Ks[:2].to(dev), W, H, sh_degree=2, packed=packed,
                                         rasterize_mode=mode, camera_model=model, absgrad=True,
                                         render_mode="RGB+ED", sparse_grad=False)
            (rc.sum() + ra.sum()).backward()
            n += 1
# the plain kernels behind tile_size != 16 and the first-generation forward kernel
from gsplat.b200_impl._lib import lib  # noqa: E402

for tile_size, variant in ((8, 2), (16, 1), (16, 3)):
    lib.gsb_set_option(0, variant)
    lib.gsb_set_option(1, 1 if variant == 1 else 2)
    p = [t.clone().to(dev).requires_grad_(True) for t in (means, quats, scales * 3 + 0.002, opac, shs)]
    rc, ra, meta = rasterization(*p, viewmats[:2].to(dev), Ks[:2].to(dev), W, H, sh_degree=2, packed=False,
                                 tile_size=tile_size, absgrad=True)
    (rc.sum() + ra.sum()).backward()
    n += 1
lib.gsb_set_option(0, 2)
lib.gsb_set_option(1, 2)
torch.cuda.synchronize()
print("sanitize_small: %d fwd+bwd passes, last n_isects=%d" % (n, meta["flatten_ids"].numel()))
