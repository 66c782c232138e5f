"""Deterministic synthetic scenes of TUM / OAK-D shape.

Scenes are generated on the CPU from seeded generators so that the CPU oracle
and the CUDA kernels see identical bits; callers move them to the GPU.
Intrinsics follow the reference's dataset tables (gslam/data.py:24, :39-43).
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import torch

TUM_FR1 = (517.3, 516.5, 318.6, 255.3, 640, 480)                 # gslam/data.py:24
OAKD = (2285.44 / 3, 2283.46 / 3, 1939.57 / 3, 1084.25 / 3, 1280, 720)   # gslam/data.py:39-43, x1/3


@dataclass
class Scene:
    means: torch.Tensor             # [N,3]
    quats: torch.Tensor             # [N,4]  un-normalised, as gslam/insertion.py:240
    log_scales: torch.Tensor        # [N,3]
    logit_opacities: torch.Tensor   # [N]
    logit_colors: torch.Tensor      # [N,3]
    log_uncertainties: torch.Tensor  # [N]
    viewmats: torch.Tensor          # [C,4,4] world->camera
    Ks: torch.Tensor                # [C,3,3]
    width: int
    height: int

    def to(self, device):
        kw = {}
        for k, v in self.__dict__.items():
            kw[k] = v.to(device) if isinstance(v, torch.Tensor) else v
        return Scene(**kw)

    def params(self):
        return (self.means, self.quats, self.log_scales, self.logit_opacities,
                self.logit_colors, self.log_uncertainties)


def intrinsics(kind: str = "tum", scale: float = 1.0):
    fx, fy, cx, cy, W, H = TUM_FR1 if kind == "tum" else OAKD
    W, H = int(round(W * scale)), int(round(H * scale))
    K = torch.tensor([[fx * scale, 0, cx * scale], [0, fy * scale, cy * scale], [0, 0, 1.0]])
    return K, W, H


def _small_se3(g: torch.Generator, max_deg: float = 2.0, max_t: float = 0.03) -> torch.Tensor:
    axis = torch.randn(3, generator=g)
    axis = axis / axis.norm()
    ang = math.radians(max_deg) * float(torch.rand(1, generator=g))
    K = torch.tensor([[0, -axis[2], axis[1]], [axis[2], 0, -axis[0]], [-axis[1], axis[0], 0]])
    R = torch.eye(3) + math.sin(ang) * K + (1 - math.cos(ang)) * (K @ K)
    T = torch.eye(4)
    T[:3, :3] = R
    T[:3, 3] = (torch.rand(3, generator=g) * 2 - 1) * max_t
    return T


def make_scene(n_gauss: int, n_views: int = 1, kind: str = "tum", scale: float = 1.0,
               seed: int = 0, pose_seed: int = 1, log_scale_mean: float = math.log(0.01),
               dtype=torch.float32) -> Scene:
    K, W, H = intrinsics(kind, scale)
    g = torch.Generator().manual_seed(seed)
    z = torch.rand(n_gauss, generator=g) * 4.5 + 0.5
    u = (torch.rand(n_gauss, generator=g) * 1.2 - 0.1) * W
    v = (torch.rand(n_gauss, generator=g) * 1.2 - 0.1) * H
    # back-project like Camera.backproject (gslam/primitives.py:388-392)
    x = (u - K[0, 2]) / K[0, 0] * z
    y = (v - K[1, 2]) / K[1, 1] * z
    means = torch.stack([x, y, z], dim=-1)
    log_scales = log_scale_mean + 0.5 * torch.randn(n_gauss, 3, generator=g)
    quats = torch.rand(n_gauss, 4, generator=g)
    opac = torch.rand(n_gauss, generator=g) * 0.75 + 0.2
    logit_opacities = torch.log(opac / (1 - opac))
    logit_colors = torch.randn(n_gauss, 3, generator=g)
    log_unc = 1.0 + 0.1 * torch.randn(n_gauss, generator=g)

    gp = torch.Generator().manual_seed(pose_seed)
    views = [torch.eye(4)]
    for _ in range(1, n_views):
        views.append(_small_se3(gp) @ views[-1])
    viewmats = torch.stack(views)
    Ks = K[None].repeat(n_views, 1, 1)
    sc = Scene(means, quats, log_scales, logit_opacities, logit_colors, log_unc, viewmats, Ks, W, H)
    if dtype != torch.float32:
        for k, val in list(sc.__dict__.items()):
            if isinstance(val, torch.Tensor):
                setattr(sc, k, val.to(dtype))
    return sc


def perturbed(scene: Scene, seed: int = 2, amount: float = 0.05) -> Scene:
    """A perturbed copy, rendered to give target images with non-trivial losses."""
    g = torch.Generator().manual_seed(seed)
    n = lambda t: t + amount * torch.randn(t.shape, generator=g).to(t.dtype)
    return Scene(scene.means + 0.01 * torch.randn(scene.means.shape, generator=g).to(scene.means.dtype),
                 scene.quats, n(scene.log_scales), n(scene.logit_opacities), n(scene.logit_colors),
                 scene.log_uncertainties, scene.viewmats, scene.Ks, scene.width, scene.height)


def target_images(n_views: int, height: int, width: int, seed: int = 2) -> torch.Tensor:
    """Smooth pseudo-random RGB targets in [0,1], [C,H,W,3] (used at sizes where the
    oracle render of a perturbed scene would take minutes)."""
    g = torch.Generator().manual_seed(seed)
    low = torch.rand(n_views, 3, max(height // 16, 2), max(width // 16, 2), generator=g)
    img = torch.nn.functional.interpolate(low, size=(height, width), mode="bilinear", align_corners=False)
    return img.permute(0, 2, 3, 1).contiguous().clamp(0, 1)


def smooth_trajectory(n_frames: int, amp_t: float = 0.25, amp_deg: float = 6.0, period: int = 240) -> torch.Tensor:
    """World->camera matrices [F,4,4] of a hand-held-like sweep that starts at identity: sinusoidal
    sideways / vertical translation (amplitude `amp_t` metres) and yaw / pitch (amplitude `amp_deg`),
    at most a few millimetres and a tenth of a degree per frame (the cfg5 sequence)."""
    out = []
    for i in range(n_frames):
        ph = 2 * math.pi * i / period
        yaw = math.radians(amp_deg) * math.sin(ph)
        pitch = math.radians(amp_deg) * 0.5 * (1 - math.cos(ph))
        Ry = torch.tensor([[math.cos(yaw), 0, math.sin(yaw)], [0, 1, 0], [-math.sin(yaw), 0, math.cos(yaw)]])
        Rx = torch.tensor([[1, 0, 0], [0, math.cos(pitch), -math.sin(pitch)], [0, math.sin(pitch), math.cos(pitch)]])
        T = torch.eye(4)
        T[:3, :3] = Rx @ Ry
        T[:3, 3] = torch.tensor([amp_t * math.sin(ph), 0.4 * amp_t * (1 - math.cos(ph)), 0.1 * amp_t * math.sin(2 * ph)])
        out.append(T)
    return torch.stack(out)
