"""Synthetic replacement of the reference's `examples/datasets/colmap.py` (Parser / Dataset), same attributes and item
layout, no COLMAP files: the "capture" is rendered once, on the GPU, from ground-truth Gaussians built on the
reference's own fixture (assets/test_garden.npz -> tests/golden/test_garden.npz: 3D points + colours + one camera).

  * ground truth: every `GT_STRIDE`-th fixture point inside [-2, 2]^3, isotropic scale = mean distance to its 3 nearest
    neighbours, random rotation, opacity in [0.6, 1], degree-0 SH from the point colour;
  * cameras: `N_VIEWS` poses on an arc around the scene (the fixture's first camera rotated about its own up axis
    through the scene centre), fixture intrinsics at half resolution (324x210);
  * images: `rasterization()` of the ground truth (whatever GSPLAT_BACKEND is active), stored as uint8;
  * "SfM points" handed to the trainer: the ground-truth centres with Gaussian noise.
"""
import os

import numpy as np
import torch

GT_STRIDE = 4
N_VIEWS = 24
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))


def _look_at(eye, target, up):
    fwd = target - eye
    fwd = fwd / np.linalg.norm(fwd)
    right = np.cross(fwd, up)  # camera x (OpenCV: x right, y down, z forward)
    right = right / np.linalg.norm(right)
    down = np.cross(fwd, right)
    c2w = np.eye(4, dtype=np.float32)
    c2w[:3, 0], c2w[:3, 1], c2w[:3, 2], c2w[:3, 3] = right, down, fwd, eye
    return c2w


def _rotate_about(p, centre, axis, angle):
    axis = axis / np.linalg.norm(axis)
    v = p - centre
    c, s = np.cos(angle), np.sin(angle)
    return centre + v * c + np.cross(axis, v) * s + axis * np.dot(axis, v) * (1 - c)


class Parser:
    def __init__(self, data_dir="", factor=1, normalize=False, test_every=8):
        from sklearn.neighbors import NearestNeighbors

        from gsplat.rendering import rasterization

        data = np.load(os.path.join(ROOT, "tests", "golden", "test_garden.npz"))
        pts = data["means3d"].astype(np.float32)
        rgb = data["colors"].astype(np.float32) / 255.0
        keep = np.all((pts >= -2) & (pts <= 2), axis=-1)
        pts, rgb = pts[keep][::GT_STRIDE], rgb[keep][::GT_STRIDE]
        rng = np.random.default_rng(0)
        dist, _ = NearestNeighbors(n_neighbors=4).fit(pts).kneighbors(pts)
        scale = np.clip(dist[:, 1:].mean(-1), 1e-3, 0.05).astype(np.float32)

        self.factor, self.normalize, self.test_every = factor, normalize, test_every
        W, H = int(data["width"]) // 2, int(data["height"]) // 2
        K = data["Ks"][0].astype(np.float32).copy()
        K[:2] *= 0.5
        c2w0 = np.linalg.inv(data["viewmats"][0].astype(np.float64))
        eye0, fwd0, up0 = c2w0[:3, 3], c2w0[:3, 2], -c2w0[:3, 1]
        depth = np.median((pts - eye0) @ fwd0)
        centre = eye0 + fwd0 * depth
        self.camtoworlds = np.stack([
            _look_at(_rotate_about(eye0, centre, up0, np.deg2rad(6.0 * (k - N_VIEWS // 2))).astype(np.float32),
                     centre.astype(np.float32), up0.astype(np.float32)) for k in range(N_VIEWS)])
        self.image_names = [f"synthetic_{k:03d}" for k in range(N_VIEWS)]
        self.image_paths = list(self.image_names)
        self.camera_ids = [0] * N_VIEWS
        self.Ks_dict = {0: K}
        self.params_dict = {0: np.empty(0, dtype=np.float32)}
        self.imsize_dict = {0: (W, H)}
        self.mask_dict = {0: None}
        self.points = (pts + rng.normal(0, 0.01, pts.shape)).astype(np.float32)
        self.points_err = np.zeros(len(pts), dtype=np.float32)
        self.points_rgb = (rgb * 255).astype(np.uint8)
        self.point_indices = {}
        self.transform = np.eye(4)
        self.bounds = np.array([0.01, 1.0])
        self.extconf = {"spiral_radius_scale": 1.0, "no_factor_suffix": False}
        cam_pos = self.camtoworlds[:, :3, 3]
        self.scene_scale = float(np.max(np.linalg.norm(cam_pos - cam_pos.mean(0), axis=1)))

        # render the capture from the ground truth
        dev = torch.device("cuda", torch.cuda.current_device())
        n = len(pts)
        quats = torch.from_numpy(rng.normal(size=(n, 4)).astype(np.float32)).to(dev)
        scales = torch.from_numpy(np.repeat(scale[:, None], 3, 1)).to(dev)
        opac = torch.from_numpy(rng.uniform(0.6, 1.0, n).astype(np.float32)).to(dev)
        sh0 = ((torch.from_numpy(rgb) - 0.5) / 0.28209479177387814)[:, None, :].to(dev)
        viewmats = torch.linalg.inv(torch.from_numpy(self.camtoworlds)).to(dev)
        Ks = torch.from_numpy(K)[None].expand(N_VIEWS, 3, 3).contiguous().to(dev)
        images = []
        with torch.no_grad():
            for lo in range(0, N_VIEWS, 8):
                rc, _, _ = rasterization(torch.from_numpy(pts).to(dev), quats, scales, opac, sh0, viewmats[lo:lo + 8],
                                         Ks[lo:lo + 8], W, H, sh_degree=0, packed=False)
                images.append((rc.clamp(0, 1) * 255).round().to(torch.uint8).cpu().numpy())
        self.images = np.concatenate(images)
        self.ground_truth_count = n


class Dataset:
    """Item layout of the reference's Dataset (datasets/colmap.py): K, camtoworld, image [H,W,3] float 0..255, image_id."""

    def __init__(self, parser, split="train", patch_size=None, load_depths=False):
        assert patch_size is None and not load_depths, "not needed by this run"
        self.parser = parser
        idx = np.arange(len(parser.image_names))
        self.indices = idx[idx % parser.test_every != 0] if split == "train" else idx[idx % parser.test_every == 0]

    def __len__(self):
        return len(self.indices)

    def __getitem__(self, item):
        i = int(self.indices[item])
        p = self.parser
        return {"K": torch.from_numpy(p.Ks_dict[0].copy()), "camtoworld": torch.from_numpy(p.camtoworlds[i].copy()),
                "image": torch.from_numpy(p.images[i].astype(np.float32)), "image_id": item}
