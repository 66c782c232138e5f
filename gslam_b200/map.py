"""The map container whose `forward` / `render` are the callers of the rasterizer.

Mirrors reference gslam/map.py:13-159 (`GaussianSplattingData`): same parameter names, constructor
order, `render(cameras, viewmats, visibility_min_T)`, `forward(cameras, poses, render_depth,
visibility_min_T)`, `empty / clone / mask / no_grad_clone / as_dict`.  The only differences are
below the call surface: `Ks` and the background are batched once per call, and the window's poses go
through one pose kernel (`primitives.pose_batch`) instead of one autograd chain per view.
"""
from __future__ import annotations

from typing import List

import torch

from .primitives import Camera, PoseZhou, pose_batch
from .rasterization import RasterizationOutput, rasterization


def create_batch(things, getter=None) -> torch.Tensor:
    """gslam/utils.py:17-23."""
    if getter is not None:
        things = [getter(t) for t in things]
    return torch.stack(things, dim=0)


class GaussianSplattingData(torch.nn.Module):
    _per_splat_params = ["means", "quats", "scales", "opacities", "colors", "log_uncertainties", "ages"]

    def __init__(self, means, quats, scales, opacities, colors, log_uncertainties, ages):
        super().__init__()
        self.means = torch.nn.Parameter(means)
        self.quats = torch.nn.Parameter(quats)
        self.scales = torch.nn.Parameter(scales)                    # log-scales
        self.opacities = torch.nn.Parameter(opacities)              # logits
        self.colors = torch.nn.Parameter(colors)                    # logits
        self.log_uncertainties = torch.nn.Parameter(log_uncertainties)
        self.ages = torch.nn.Parameter(ages, requires_grad=False)
        self.register_buffer("background", torch.tensor([0.0, 0.0, 0.0], device=self.means.device).float())

    def _rasterize(self, cameras: List[Camera], viewmats: torch.Tensor, render_mode: str,
                   visibility_min_T: float) -> RasterizationOutput:
        Ks = create_batch(cameras, lambda c: c.intrinsics)
        return rasterization(
            means=self.means, quats=self.quats, log_scales=self.scales, logit_opacities=self.opacities,
            logit_colors=self.colors, viewmats=viewmats, Ks=Ks, width=cameras[0].width,
            height=cameras[0].height, render_mode=render_mode, packed=False,
            log_uncertainties=self.log_uncertainties, visibility_min_T=visibility_min_T,
            backgrounds=self.background.tile([len(cameras), 1]))

    def render(self, cameras: List[Camera], viewmats: List[torch.Tensor], visibility_min_T: float = 0.5):
        """gslam/map.py:50-76: render given viewmats (no pose modules), always RGB+D."""
        return self._rasterize(cameras, create_batch(viewmats), "RGB+D", visibility_min_T)

    def forward(self, cameras: List[Camera], poses: List[PoseZhou], render_depth: bool = False,
                visibility_min_T: float = 0.5) -> RasterizationOutput:
        """gslam/map.py:78-103."""
        return self._rasterize(cameras, pose_batch(poses), "RGB+D" if render_depth else "RGB", visibility_min_T)

    @staticmethod
    def empty(device: str = "cuda") -> "GaussianSplattingData":
        e = lambda: torch.tensor([], device=device)  # noqa: E731
        return GaussianSplattingData(e(), e(), e(), e(), e(), e(), e().long())

    def clone(self) -> "GaussianSplattingData":
        return GaussianSplattingData(*(getattr(self, p).clone().detach() for p in self._per_splat_params[:-1]),
                                     self.ages.clone())

    def mask(self, m) -> "GaussianSplattingData":
        return GaussianSplattingData(self.means[m, :], self.quats[m, :], self.scales[m, :], self.opacities[m],
                                     self.colors[m, :], self.log_uncertainties[m], self.ages[m])

    def no_grad_clone(self) -> "GaussianSplattingData":
        ret = self.clone()
        for p in self._per_splat_params:
            getattr(ret, p).requires_grad_(False)
        return ret

    def as_dict(self):
        return torch.nn.ParameterDict({p: getattr(self, p) for p in self._per_splat_params})
