"""Synthetic scenes of BASELINE.json configs[2]/[3], shared by the tools."""
import math

import torch


def ring_cameras(C, W, H, radius=6.0):
    """C look-at cameras on a ring of `radius` around the origin (world -> camera matrices), fov 60 degrees."""
    f = W / (2 * math.tan(math.radians(30)))
    vms = []
    for i in range(C):
        az = 2 * math.pi * i / C
        eye = torch.tensor([radius * math.cos(az), 0.0, radius * math.sin(az)])
        fwd = -eye / eye.norm()
        right = torch.linalg.cross(torch.tensor([0.0, 1.0, 0.0]), fwd)
        right = right / right.norm()
        up = torch.linalg.cross(fwd, right)
        R = torch.stack([right, up, fwd])
        vm = torch.eye(4)
        vm[:3, :3], vm[:3, 3] = R, -R @ eye
        vms.append(vm)
    Ks = torch.tensor([[f, 0, W / 2], [0, f, H / 2], [0, 0, 1.0]]).expand(C, 3, 3).contiguous()
    return torch.stack(vms), Ks


def ring_scene(N, C, W, H, seed=0):
    """means U[-3,3]^3, scales exp(U[ln .002, ln .02]), SH degree 3 coefficients; CPU tensors."""
    g = torch.Generator().manual_seed(seed)
    means = torch.rand(N, 3, generator=g) * 6 - 3
    quats = torch.randn(N, 4, generator=g)
    lo, hi = math.log(0.002), math.log(0.02)
    scales = torch.exp(torch.rand(N, 3, generator=g) * (hi - lo) + lo)
    opac = torch.rand(N, generator=g)
    shs = torch.randn(N, 16, 3, generator=g) * 0.3
    shs[:, 0] += 0.5
    viewmats, Ks = ring_cameras(C, W, H)
    return means, quats, scales, opac, shs, viewmats, Ks
