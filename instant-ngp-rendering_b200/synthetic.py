"""Synthetic NeRF datasets for tests and bench.py (there is no network for real datasets and the
reference's data directory does not travel to the GPU box): an analytic emissive volume of
Gaussian blobs, rendered with torch into posed RGBA8 images of any shape (e.g. the fox shape,
50 x 1080 x 1920, aabb_scale 4). Data generation only; never part of a timed region.
"""
import math

import numpy as np
import torch


def look_at_ngp(eye, target=(0.5, 0.5, 0.5), up=(0.0, 0.0, 1.0)):
    """Camera-to-world 3x4 in the NGP convention: columns right, down, forward, origin."""
    eye = np.asarray(eye, dtype=np.float64)
    fwd = np.asarray(target, dtype=np.float64) - eye
    fwd /= np.linalg.norm(fwd)
    right = np.cross(fwd, np.asarray(up, dtype=np.float64))
    right /= np.linalg.norm(right)
    down = np.cross(fwd, right)
    m = np.stack([right, down, fwd, eye], axis=1)
    return np.ascontiguousarray(m, dtype=np.float32)


def orbit_poses(n, radius=1.1, center=(0.5, 0.5, 0.5), seed=0):
    """n cameras on a sphere around `center`, spread with a golden-angle spiral on the upper 3/4."""
    poses = []
    for i in range(n):
        z = 0.75 - 1.25 * (i + 0.5) / n          # from above to slightly below
        phi = i * math.pi * (3.0 - math.sqrt(5.0))
        r = math.sqrt(max(1.0 - z * z, 0.0))
        eye = np.asarray(center) + radius * np.array([r * math.cos(phi), r * math.sin(phi), z])
        poses.append(look_at_ngp(eye, center))
    return poses


class BlobScene:
    """sigma(x) = sum_k a_k exp(-|x - c_k|^2 / (2 s_k^2)); colour = density-weighted mix of blob colours."""

    def __init__(self, n_blobs=6, seed=7, device="cuda", amp_scale=1.0):
        g = np.random.default_rng(seed)
        self.centers = torch.tensor(0.5 + (g.random((n_blobs, 3)) - 0.5) * 0.5, dtype=torch.float32, device=device)
        self.sigmas = torch.tensor(0.04 + g.random(n_blobs) * 0.06, dtype=torch.float32, device=device)
        # amp_scale > 1 sharpens the blobs towards solid objects (earlier ray termination)
        self.amps = torch.tensor((30.0 + g.random(n_blobs) * 60.0) * amp_scale, dtype=torch.float32, device=device)
        self.colors = torch.tensor(0.15 + 0.85 * g.random((n_blobs, 3)), dtype=torch.float32, device=device)
        self.device = device

    def field(self, x):
        d2 = ((x[:, None, :] - self.centers[None]) ** 2).sum(-1)
        w = self.amps[None] * torch.exp(-d2 / (2 * self.sigmas[None] ** 2))
        sigma = w.sum(-1)
        rgb = (w[:, :, None] * self.colors[None]).sum(1) / (sigma[:, None] + 1e-8)
        return sigma, rgb

    @torch.no_grad()
    def render(self, pose, width, height, focal, n_steps=96, near=0.2, far=2.0, chunk=1 << 19):
        """Linear premultiplied RGBA float32 [h][w][4] by uniform quadrature along each pixel ray."""
        dev = self.device
        m = torch.tensor(pose, dtype=torch.float32, device=dev)
        ys, xs = torch.meshgrid(torch.arange(height, device=dev), torch.arange(width, device=dev), indexing="ij")
        u = (xs.reshape(-1).float() + 0.5) / width
        v = (ys.reshape(-1).float() + 0.5) / height
        out = torch.empty((width * height, 4), dtype=torch.float32, device=dev)
        ts = near + (far - near) * (torch.arange(n_steps, device=dev).float() + 0.5) / n_steps
        dt = (far - near) / n_steps
        for s in range(0, width * height, chunk):
            uu, vv = u[s:s + chunk], v[s:s + chunk]
            d = torch.stack([(uu - 0.5) * width / focal, (vv - 0.5) * height / focal, torch.ones_like(uu)], -1)
            d = d @ m[:, :3].T
            d = d / d.norm(dim=-1, keepdim=True)
            rgb_acc = torch.zeros((uu.numel(), 3), device=dev)
            T = torch.ones(uu.numel(), device=dev)
            for t in ts:
                sigma, rgb = self.field(m[:, 3][None] + d * t)
                a = 1 - torch.exp(-sigma * dt)
                rgb_acc += (T * a)[:, None] * rgb
                T = T * (1 - a)
            out[s:s + chunk, :3] = rgb_acc
            out[s:s + chunk, 3] = 1 - T
        return out.reshape(height, width, 4)


def linear_premul_to_srgb8(rgba):
    """Linear premultiplied float RGBA -> sRGB8 straight alpha (what a PNG/JPEG loader hands the reference)."""
    a = rgba[..., 3:4].clamp(0, 1)
    rgb = torch.where(a > 1e-6, rgba[..., :3] / a.clamp_min(1e-6), torch.zeros_like(rgba[..., :3])).clamp(0, 1)
    srgb = torch.where(rgb < 0.0031308, 12.92 * rgb, 1.055 * rgb.clamp_min(1e-8) ** (1 / 2.4) - 0.055)
    return (torch.cat([srgb, a], -1) * 255.0 + 0.5).clamp(0, 255).to(torch.uint8)


def make_dataset(testbed, n_images=8, width=160, height=120, aabb_scale=1, focal=None, seed=7, radius=None, opaque_background=False, amp_scale=1.0, n_blobs=6):
    """Fills `testbed` with a synthetic posed dataset; returns (scene, poses, focal)."""
    scene = BlobScene(n_blobs=n_blobs, seed=seed, device=testbed._device, amp_scale=amp_scale)
    focal = focal or 1.1 * width
    radius = radius or (1.1 if aabb_scale == 1 else 1.4)
    poses = orbit_poses(n_images, radius=radius)
    testbed.create_empty_nerf_dataset(n_images, aabb_scale=aabb_scale)
    for i, p in enumerate(poses):
        img = scene.render(p, width, height, focal)
        if opaque_background:
            img[..., :3] += (1 - img[..., 3:4]) * 0.5
            img[..., 3] = 1.0
        testbed.set_nerf_image(i, linear_premul_to_srgb8(img))
        testbed.nerf.training.set_camera_extrinsics(i, p, convert_to_ngp=False)
        testbed.nerf.training.set_camera_intrinsics(i, fx=focal, fy=focal, cx=width / 2, cy=height / 2)
    return scene, poses, focal


NERF_BASE_CONFIG = {
    # configs/nerf/base.json of the reference (this fork: L=8, F=4)
    "loss": {"otype": "Huber"},
    "optimizer": {"otype": "Ema", "decay": 0.95, "nested": {"otype": "ExponentialDecay", "decay_start": 20000, "decay_interval": 10000, "decay_base": 0.33,
                  "nested": {"otype": "Adam", "learning_rate": 1e-2, "beta1": 0.9, "beta2": 0.99, "epsilon": 1e-15, "l2_reg": 1e-6}}},
    "encoding": {"otype": "HashGrid", "n_levels": 8, "n_features_per_level": 4, "log2_hashmap_size": 19, "base_resolution": 16},
    "network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 1},
    "dir_encoding": {"otype": "Composite", "nested": [{"n_dims_to_encode": 3, "otype": "SphericalHarmonics", "degree": 4}, {"otype": "Identity"}]},
    "rgb_network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 2},
}
