"""Headless tracking + mapping loop over a frame sequence (BASELINE cfg5; the callers either side
of the render path).

The reference runs this as three processes joined by queues (main.py:61-94): the sensor stream,
`Frontend.run` (gslam/frontend.py) and `Backend.run` (gslam/backend.py:797-898).  Queues, process
start-up, rerun / viser logging and dataset readers are control plane and stay out of scope; what
is kept is the ORDER OF WORK on the render path, executed inline on one CUDA stream:

    frame 0      Backend.initialize (:604-631): 5000 splats from a mock depth map, then
                 optimize_map(num_iters_initialization, prune=False), sync
    frame i > 0  Frontend.track (frontend.py:170-247): constant-motion start, igs_track_lbfgs
                 Backend, ADD_FRAME (:856-873): to_insert_keyframe (:741-795) -> add_keyframe
                 (:633-679) -> optimize_map(1, prune=True, regularize=False); sync every 5th frame
                 Backend, idle (:846-852): optimize_map(), run_pruning(), optimize_poses_lbfgs() until
                 StopOnPlateau pauses it

Every step is the corresponding module of this package: `tracking.Tracker` (CUDA-graph closure),
`mapping.MappingStep` (flat [15 N] layout, fused losses, one-launch Adam), `insertion.
InsertFromDepthMap`, `pruning.prune_round`, `surgery` (one-launch append / compaction).  The map
hand-off of `Backend.sync` -> `Frontend.sync` (deepcopy + pickle through an mp.Queue in the
reference, backend.py:508-519) is one device-to-device copy of the flat parameter buffer
(`Tracker.set_map`).

Not reproduced: the pose-graph / loop-closure bookkeeping (`enable_pgo`, off by default).
"""
from __future__ import annotations

import math
import time
from dataclasses import dataclass, field
from typing import Dict, List, Optional

import numpy as np
import torch
from torch import Tensor

from .lbfgs import lbfgs_minimize
from .losses import photometric_ssim_loss
from .graph import GraphedStep, shared_pool
from .insertion import InsertFromDepthMap, InsertUsingImagePlaneGradients, KeyframeView
from .map import GaussianSplattingData
from .mapping import (DEFAULT_LRS, LOCAL, PARAM_NAMES, FlatSplats, MappingStep, _world, allreduce_sum_,
                      gather_adam_state)
from .optim import FusedAdam
from .primitives import Camera, PoseZhou, pose_batch
from .pruning import prune_round
from .rasterization import RasterizationOutput, rasterization
from .tracking import StopOnPlateau, Tracker, TrackingConfig, constant_motion, affine_inverse


@dataclass
class MapConfig:
    """The fields of the reference MapConfig (gslam/backend.py:43-113) this loop reads, same defaults."""
    isotropic_regularization_weight: float = 0.0005
    depth_regularization_weight: float = 0.000001
    pose_optim_lr: float = 0.003
    opacity_decay: float = 0.995
    initial_opacity: float = 0.3
    initial_scale: float = 1.0
    optim_window_last_n_keyframes: int = 8
    num_iters_mapping: int = 15
    num_iters_initialization: int = 400
    opacity_pruning_threshold: float = 0.2
    size_pruning_threshold: int = 256
    ssim_weight: float = 0.2
    enable_visibility_pruning: bool = False
    kf_m: float = 0.15
    kf_cos: float = math.cos(math.pi / 30)
    n_init_splats: int = 5000               # the literal of Backend.initialize (:626)
    n_keyframe_splats: int = 100            # the literal of Backend.add_keyframe (:646)
    sync_every: int = 5                     # `frame.index % 5 == 0` (:871)
    densify_every: int = 200                # `(self.total_step % 200) == 0` (:328)
    grow_grad2d: float = 0.0002             # InsertUsingImagePlaneGradients(0.0002, 0.01) (:147-150)
    grow_scale3d: float = 0.01
    use_gt_depths: bool = False             # RGB-D mode (:111): depth terms in mapping and tracking
    graph_mapping: bool = True              # replay rounds of >= graph_min_iters iterations as a CUDA graph
    graph_min_iters: int = 10
    graph_check_every: int = 5
    optimize_poses: bool = True             # optimize_poses_lbfgs after every idle mapping round (:851)
    deterministic: bool = False             # fixed-order gradient accumulation in mapping AND tracking
                                            # (ops.deterministic): the same sequence gives the same keyframes,
                                            # map and trajectory, bit for bit, from run to run


@dataclass
class Keyframe:
    index: int
    img: Tensor                 # [H,W,3]
    pose: PoseZhou
    exposure: Tensor            # [2], fixed once the frame is a keyframe (:660-662)
    est_depths: Optional[Tensor] = None
    gt_depth: Optional[Tensor] = None


@dataclass
class _Window:
    """The static tensors of one optimize_map round."""
    poses: List[PoseZhou]
    gt: Tensor
    exposure: Tensor
    Ks: Tensor
    gt_depths: Optional[Tensor]


@dataclass
class _MappingGraph:
    g: GraphedStep
    res: object                  # MappingStepResult whose tensors every replay rewrites
    ring: Tensor                 # photometric loss of replay i at ring[i % 64]
    pos: Tensor
    pose_grads: list             # [(parameter, graph-owned gradient tensor)]
    count: int = 0
    snap: Optional[list] = None  # copies of everything a replay mutates, taken before each chunk (rollback on overflow)


@dataclass
class SlamStats:
    frames: int = 0
    keyframes: int = 0
    tracking_evals: int = 0
    tracking_syncs: int = 0
    mapping_iters: int = 0
    inserted: int = 0
    pruned: int = 0
    densified: int = 0
    pose_lbfgs_evals: int = 0
    pose_lbfgs_overflows: int = 0
    mapping_overflow_rollbacks: int = 0
    t_tracking: float = 0.0
    t_mapping: float = 0.0
    n_splats: int = 0
    last_tracking_loss: float = float("nan")
    last_photometric: float = float("nan")
    first_photometric: float = float("nan")
    history: List[Dict] = field(default_factory=list)


class Slam:
    def __init__(self, camera: Camera, map_conf: Optional[MapConfig] = None,
                 track_conf: Optional[TrackingConfig] = None, device="cuda", seed: int = 0, group=LOCAL):
        """group: LOCAL (default) maps on this GPU alone.  A process group (None = WORLD) shards the MAPPING
        window by view over its ranks (mapping.MappingStep): every rank runs this same loop on the same
        frames -- tracking, keyframe decisions, insertion and the window-pose L-BFGS are replicated, the
        mapping iterations are sharded -- which needs MapConfig(deterministic=True), so that the replicated
        arithmetic is bit-identical and every rank takes the same decisions (and issues the same collectives)."""
        self.conf = map_conf or MapConfig()
        self.group = group
        self.rank, self.world = _world(group)
        if self.world > 1 and not self.conf.deterministic:
            raise ValueError("a view-sharded Slam needs MapConfig(deterministic=True)")
        self.camera = camera.to(device)
        self.dev = torch.device(device)
        if self.dev.type != "cuda":
            raise RuntimeError("gslam_b200.slam.Slam needs a CUDA device (no CPU fallback)")
        self.gen = torch.Generator(device=self.dev).manual_seed(seed)
        track_conf = track_conf or TrackingConfig()
        track_conf.use_gt_depths = self.conf.use_gt_depths
        track_conf.deterministic = track_conf.deterministic or self.conf.deterministic
        self.tracker = Tracker(self.camera, track_conf, device)
        c = self.conf
        self.insertion = InsertFromDepthMap(0.1 * c.initial_scale, 0.2 * c.initial_scale, 0.1,
                                            c.initial_opacity, False)             # backend.py:139-146
        self.insertion_3dgs = InsertUsingImagePlaneGradients(c.grow_grad2d, c.grow_scale3d)
        self.keyframes: Dict[int, Keyframe] = {}
        self.frame_poses: List[Tensor] = []          # tracked world->camera matrix of every frame
        self.last_exposure: Optional[Tensor] = None
        self.pause_map_optim = False
        self.total_step = 0
        self.stats = SlamStats()
        empty = [torch.zeros((0, w) if w > 1 else (0,), device=self.dev) for w in (3, 4, 3, 1, 3, 1)]
        self.splats = FlatSplats(empty, align=self.world)
        # one optimizer, one group per parameter: the reference's six Adam instances (:565-590)
        self.opt = FusedAdam([{"params": [p], "lr": DEFAULT_LRS[name]}
                              for p, name in zip(self.splats.params, PARAM_NAMES)], capturable=True)
        self.pose_opt: Optional[FusedAdam] = None
        self.step = MappingStep(self.splats, self.opt, ssim_weight=c.ssim_weight,
                                isotropic_weight=c.isotropic_regularization_weight,
                                opacity_decay=c.opacity_decay,
                                group=group, comm="nccl")    # LOCAL: no collective, whatever the job size
        self.bg1 = torch.zeros(1, 3, device=self.dev)

    # ---- rendering helpers ----------------------------------------------------------------------
    @torch.no_grad()
    def render(self, viewmats: Tensor) -> RasterizationOutput:
        P = self.splats.params
        C = viewmats.shape[0]
        Ks = self.camera.intrinsics.expand(C, 3, 3).contiguous()
        return rasterization(P[0], P[1], P[2], P[3], P[4], viewmats.contiguous(), Ks, self.camera.width,
                             self.camera.height, packed=False, render_mode="RGB+D", log_uncertainties=P[5],
                             backgrounds=self.bg1.expand(C, 3).contiguous())

    def _before_surgery(self) -> None:
        """A view-sharded step keeps a Gaussian's Adam moments current only on the rank that owns its range; insertion,
        densification and pruning re-partition the Gaussians, so every rank first collects all moments."""
        if self.world > 1:
            gather_adam_state(self.splats, self.opt, self.group)

    def _pose_batch(self, poses) -> Tensor:
        return pose_batch(poses) if poses else torch.zeros(0, 4, 4, device=self.dev)

    def _views(self) -> List[KeyframeView]:
        with torch.no_grad():
            return [KeyframeView(self.camera.intrinsics, kf.pose(), kf.est_depths)
                    for kf in self.keyframes.values() if kf.est_depths is not None]

    # ---- backend ---------------------------------------------------------------------------------
    def initialize(self, img: Tensor, gt_depth: Optional[Tensor] = None) -> None:
        """Backend.initialize (gslam/backend.py:604-631)."""
        H, W = self.camera.height, self.camera.width
        d = self.dev
        mock_depth = torch.ones(1, H, W, device=d)
        mock_depth = (mock_depth + (torch.randn(1, H, W, device=d, generator=self.gen) - 0.5) * 0.3) \
            * self.conf.initial_scale
        mock_alphas = torch.full((1, H, W, 1), 0.01, device=d)
        mock = RasterizationOutput(None, mock_alphas, mock_depth)
        pose = PoseZhou(torch.eye(4, device=d)).to(d)
        kf = Keyframe(0, img, pose, torch.zeros(2, device=d), gt_depth=gt_depth)
        self.keyframes[0] = kf
        self._before_surgery()
        self.stats.inserted += self.insertion.step(self.splats, self.opt, mock, self.camera.intrinsics,
                                                   torch.eye(4, device=d), img, self.conf.n_init_splats,
                                                   gt_depthmap=gt_depth if self.conf.use_gt_depths else None,
                                                   generator=self.gen)
        self.stats.keyframes = 1

    def add_keyframe(self, index: int, img: Tensor, viewmat: Tensor, exposure: Tensor,
                     gt_depth: Optional[Tensor] = None) -> None:
        """Backend.add_keyframe (gslam/backend.py:633-679)."""
        out = self.render(viewmat.view(1, 4, 4))
        out.depthmaps = out.depthmaps * self.conf.initial_scale
        self._before_surgery()
        self.stats.inserted += self.insertion.step(self.splats, self.opt, out, self.camera.intrinsics, viewmat,
                                                   img, self.conf.n_keyframe_splats, keyframes=self._views(),
                                                   generator=self.gen)
        pose = PoseZhou(viewmat.detach().clone()).to(self.dev)
        self.keyframes[index] = Keyframe(index, img, pose, exposure.detach().clone(), out.depthmaps[0].clone(),
                                         gt_depth)
        group = {"params": [pose.dR, pose.dt], "lr": self.conf.pose_optim_lr}
        if self.pose_opt is None:
            self.pose_opt = FusedAdam([group])
        else:
            self.pose_opt.add_param_group(group)
        self.stats.keyframes = len(self.keyframes)

    @torch.no_grad()
    def to_insert_keyframe(self, viewmat: Tensor) -> bool:
        """Backend.to_insert_keyframe (gslam/backend.py:741-795): translation against the median
        rendered depth, then the angle between the optical axes."""
        prev = self.keyframes[max(self.keyframes)].pose()
        out = self.render(torch.stack([viewmat, prev]))
        diff = affine_inverse(viewmat) @ prev
        translation = diff[:3, 3].norm()
        seen = out.depthmaps[out.alphas[..., 0] > 0.1]
        median_depth = seen.median() if seen.numel() else torch.ones((), device=self.dev)
        cos = torch.nn.functional.cosine_similarity(viewmat[:3, 2], prev[:3, 2], dim=0)
        decide = torch.stack([translation > self.conf.kf_m * median_depth, cos < self.conf.kf_cos])
        return bool(decide.any().item())                # ONE host read for both criteria

    def optimization_window(self) -> List[Keyframe]:
        """Backend.optimization_window (gslam/backend.py:197-246), enable_pgo=False branch: the last n
        keyframes (the reference's `min(0, ...)` makes the random part empty)."""
        keys = sorted(self.keyframes)[-self.conf.optim_window_last_n_keyframes:]
        return [self.keyframes[k] for k in keys]

    def _mapping_iteration(self, w: "_Window", densify: bool):
        """One iteration of the optimize_map loop body (gslam/backend.py:258-359), eagerly."""
        c = self.conf
        for p in w.poses:
            p.dR.grad = None
            p.dt.grad = None
        self.step.retain_means2d_grad = densify
        res = self.step(self._pose_batch(w.poses), w.Ks, w.gt, w.exposure, self.camera.width, self.camera.height,
                        self._n_window, step_optimizer=not densify, gt_depths=w.gt_depths)
        changed = False
        if densify:
            # the reference densifies between backward and the optimizer step (:328-351): the
            # parameters it re-creates carry no gradient, so that iteration takes no splat step and
            # only the opacity decay applies (radii padded with zeros for the new rows)
            self._before_surgery()
            n_dup, n_split = self.insertion_3dgs.step(self.splats, self.opt, res.outputs, self.gen,
                                                      group=self.group, n_views_total=self._n_window)
            self.stats.inserted += n_dup + n_split
            self.stats.densified += n_dup + n_split
            changed = n_dup + n_split > 0
            if changed:
                with torch.no_grad():
                    seen_twice = torch.cat([res.vis_count > 1,          # reduced over the window's views (all ranks)
                                            res.vis_count.new_zeros(n_dup + n_split, dtype=torch.bool)])
                    self.splats.params[3].mul_(torch.where(seen_twice, c.opacity_decay, 1.0))
                    res.radii = res.outputs.radii
                    res.n_touched = torch.cat([res.n_touched, res.n_touched.new_zeros(
                        res.n_touched.shape[0], n_dup + n_split)], dim=1)
            else:
                self.step.apply_pending()        # nothing was re-created: the iteration's update after all
        if self.pose_opt is not None:
            self.pose_opt.step()            # keyframe 0 is in no group: its pose stays fixed (:604-611)
        return res, changed

    @staticmethod
    def _light(res):
        """What the end of a round reads from an iteration's result, detached: holding the live
        RasterizationOutput would keep its autograd graph -- and the AccumulateGrad nodes of the
        parameters, which remember the stream they were created on -- alive into the next capture
        (cudaErrorStreamCaptureImplicit)."""
        from types import SimpleNamespace
        from .mapping import MappingStepResult
        return MappingStepResult(res.total.detach(), res.photometric.detach(), res.radii, res.n_touched,
                                 res.vis_count, SimpleNamespace(depthmaps=res.outputs.depthmaps.detach()))

    def _capture_mapping(self, w: "_Window") -> "_MappingGraph":
        """The iteration above (without densification) as one CUDA graph over the window's static
        tensors; the photometric loss of every replay goes to a device ring, the pose gradients stay
        in graph-owned buffers that the (eager, per-keyframe step count) pose optimizer reads."""
        from . import ops
        ring = torch.zeros(64, device=self.dev)
        pos = torch.zeros((), dtype=torch.long, device=self.dev)
        box = {}

        def fn():
            for p in w.poses:
                p.dR.grad = None
                p.dt.grad = None
            self.step.retain_means2d_grad = False
            res = self.step(self._pose_batch(w.poses), w.Ks, w.gt, w.exposure, self.camera.width, self.camera.height,
                            self._n_window, gt_depths=w.gt_depths)
            with torch.no_grad():
                ring.index_copy_(0, pos.view(1) % ring.shape[0], res.photometric.detach().view(1))
                pos.add_(1)
            box["res"] = self._light(res)
            return box["res"].total

        hint = int(ops.last_n_isects.item()) if (w.poses and ops.last_n_isects is not None) else 1
        g = GraphedStep(fn, headroom=1.5, n_isects_hint=hint, pool=shared_pool(self.dev))
        grads = [(q, q.grad) for p in w.poses for q in (p.dR, p.dt)]
        return _MappingGraph(g, box["res"], ring, pos, grads)

    def _replay_state(self, w: "_Window") -> list:
        """Every tensor a replayed mapping iteration changes: the flat parameter buffer, the Adam moments and
        device step counter of the splats, the window's poses and their optimizer state."""
        ts = [self.splats.flat]
        for st in self.opt.state.values():
            ts += [st["exp_avg"], st["exp_avg_sq"]]
        if self.opt._dev_step is not None:
            ts.append(self.opt._dev_step)
        for p in w.poses:
            for q in (p.dR, p.dt):
                ts.append(q.data)
                st = self.pose_opt.state.get(q) if self.pose_opt is not None else None
                if st:
                    ts += [st["step"], st["exp_avg"], st["exp_avg_sq"]]
        return ts

    @staticmethod
    def _copy_all(dst: list, src: list) -> None:
        on_dev = [(d, s_) for d, s_ in zip(dst, src) if d.is_cuda]
        if on_dev:
            torch._foreach_copy_([d for d, _ in on_dev], [s_ for _, s_ in on_dev])
        for d, s_ in zip(dst, src):
            if not d.is_cuda:
                d.copy_(s_)

    def optimize_map(self, n_iters: Optional[int] = None, prune: bool = True, regularize: bool = True) -> None:
        """Backend.optimize_map (gslam/backend.py:248-409).

        Rounds of `graph_min_iters` or more iterations are replayed as a CUDA graph after the first
        (eager) one: the map size and the window are fixed inside a round, so the ~50 launches of an
        iteration and its two host syncs (intersection count, loss read-back) collapse to one launch.
        The reference reads the loss every iteration for StopOnPlateau(3, 0.012); the replayed
        iterations are checked in chunks of `graph_check_every`, so a stop can come that many minus
        one iterations late.  Densification iterations (every 200th) always run eagerly."""
        c = self.conf
        n_iters = c.num_iters_mapping if n_iters is None else n_iters
        stopper = StopOnPlateau(3, 0.012)
        self.step.tv_w = c.depth_regularization_weight if regularize else 0.0
        window_all = self.optimization_window()
        # a view belongs to the rank given by its keyframe's position in the (append-only) keyframe list, so
        # its owner -- and the Adam state of its pose -- never changes while the window slides
        pos = {k: i for i, k in enumerate(sorted(self.keyframes))}
        window = [k for k in window_all if pos[k.index] % self.world == self.rank]
        self._n_window = len(window_all)
        H, W = self.camera.height, self.camera.width
        w = _Window([k.pose for k in window],
                    torch.stack([k.img for k in window]) if window else torch.zeros(0, H, W, 3, device=self.dev),
                    torch.stack([k.exposure for k in window]) if window else torch.zeros(0, 2, device=self.dev),
                    self.camera.intrinsics.expand(len(window), 3, 3).contiguous(),
                    (torch.stack([k.gt_depth for k in window]) if window else torch.zeros(0, H, W, device=self.dev))
                    if c.use_gt_depths else None)
        for k in self.keyframes.values():            # pose_optimizer.zero_grad() (:265): nothing stale steps
            k.pose.dR.grad = None
            k.pose.dt.grad = None
        use_graph = c.graph_mapping and n_iters >= c.graph_min_iters
        graph: Optional[_MappingGraph] = None
        res, done, stopped = None, 0, False

        def account(photo: float) -> bool:
            self.stats.mapping_iters += 1
            self.stats.last_photometric = photo
            if self.stats.mapping_iters == 1:
                self.stats.first_photometric = photo
            return stopper.stop(photo)

        def until_densify() -> int:
            if c.densify_every <= 0:
                return 1 << 30
            return c.densify_every - 1 - (self.total_step % c.densify_every)

        while done < n_iters and not stopped:
            chunk = min(c.graph_check_every, n_iters - done, until_densify()) if use_graph and done > 0 else 0
            if chunk <= 0:
                self.total_step += 1
                densify = c.densify_every > 0 and self.total_step % c.densify_every == 0
                res, changed = self._mapping_iteration(w, densify)
                res = self._light(res)
                if densify:
                    prune = False
                    if changed:
                        graph = None
                done += 1
                stopped = account(float(res.photometric))      # the reference's per-iteration .item() (:338-349)
                continue
            if graph is None:
                graph = self._capture_mapping(w)
            start = graph.count
            # a replay that overflows its captured key buffers renders a truncated image, and the chunk's updates
            # are applied before the flag is read: keep a copy of everything the chunk mutates and roll back
            live = self._replay_state(w)
            if graph.snap is None or len(graph.snap) != len(live):
                graph.snap = [torch.empty_like(t) for t in live]
            with torch.no_grad():
                self._copy_all(graph.snap, live)
            prev_res = res
            for _ in range(chunk):
                graph.g.replay()
                if self.pose_opt is not None:
                    for q, gr in graph.pose_grads:
                        q.grad = gr
                    self.pose_opt.step()
            graph.count += chunk
            self.total_step += chunk
            done += chunk
            res = graph.res
            losses = graph.ring.tolist()                        # the chunk's one sync
            overflowed = graph.g.overflow
            if self.world > 1:
                # a rank that redoes a chunk issues the chunk's collectives again: every rank must take the same
                # branch, so the flag is the MAX over the ranks (one 4-byte all-reduce per chunk)
                from .mapping import allreduce_max_
                overflowed = allreduce_max_(overflowed.clone(), self.group)
            if int(overflowed.item()) != 0:
                # more intersections than the captured buffers hold: the replayed iterations rendered
                # a truncated image.  Undo the chunk (parameters, moments, step counters, poses), do not
                # account its losses, and redo it eagerly (exact buffers) from here on
                with torch.no_grad():
                    self._copy_all(live, graph.snap)
                self.total_step -= chunk
                done -= chunk
                res = prev_res
                use_graph, graph = False, None
                self.stats.mapping_overflow_rollbacks += 1
                continue
            for i in range(start, start + chunk):
                if account(losses[i % len(losses)]):
                    stopped = True
                    break
        if stopped:
            self.pause_map_optim = True
        for k in self.keyframes.values():
            k.pose.dR.grad = None
            k.pose.dt.grad = None
        if res is None:
            return
        with torch.no_grad():
            for k, dmap in zip(window, res.outputs.depthmaps):
                k.est_depths = dmap.detach().clone()
            if self.world > 1:
                self._share_window(window_all, pos)
            if prune:
                self._before_surgery()
                self.stats.pruned += prune_round(
                    self.splats, self.opt, res.radii, res.n_touched, c.opacity_pruning_threshold,
                    c.size_pruning_threshold, 3 if c.enable_visibility_pruning else None,
                    c.optim_window_last_n_keyframes, group=self.group,
                    n_views_total=len(window_all) if self.world > 1 else None)
        self.stats.n_splats = self.splats.n

    @torch.no_grad()
    def _share_window(self, window_all, pos) -> None:
        """End of a sharded round: every rank gets the poses and the estimated depth maps of the views the
        other ranks optimised (9 floats per view + one depth map per view, one all-reduce each:
        the owner contributes the value, everybody else zeros)."""
        H, W = self.camera.height, self.camera.width
        poses = torch.zeros(len(window_all), 9, device=self.dev)
        depths = torch.zeros(len(window_all), H, W, device=self.dev)
        for i, k in enumerate(window_all):
            if pos[k.index] % self.world == self.rank:
                poses[i, :6], poses[i, 6:] = k.pose.dR.detach(), k.pose.dt.detach()
                if k.est_depths is not None:
                    depths[i] = k.est_depths
        allreduce_sum_(poses, self.group)
        allreduce_sum_(depths, self.group)
        for i, k in enumerate(window_all):
            if pos[k.index] % self.world != self.rank:
                k.pose.dR.copy_(poses[i, :6])
                k.pose.dt.copy_(poses[i, 6:])
                k.est_depths = depths[i].clone()

    @torch.no_grad()
    def run_pruning(self) -> None:
        """Backend.run_pruning (gslam/backend.py:411-434): size and opacity criteria seen from the
        latest keyframe."""
        last = self.keyframes[max(self.keyframes)]
        out = self.render(last.pose().view(1, 4, 4))
        c = self.conf
        self._before_surgery()
        self.stats.pruned += prune_round(self.splats, self.opt, out.radii, out.n_touched,
                                         c.opacity_pruning_threshold, c.size_pruning_threshold, None,
                                         c.optim_window_last_n_keyframes, group=LOCAL)   # replicated: no collective
        self.stats.n_splats = self.splats.n

    def optimize_poses_lbfgs(self) -> Optional[float]:
        """Backend.optimize_poses_lbfgs (gslam/backend.py:436-506): L-BFGS (history 10, strong Wolfe,
        tolerance_change 1e-7, lr 1) over the window's poses except keyframe 0, objective = the
        photometric part of the mapping loss on the frozen map.  The 9 W unknowns live on the host
        (gslam_b200.lbfgs); an evaluation is one upload, one render + loss + backward to the poses
        (no gradient to the map, which the reference computes and discards) and one download --
        after the first, one CUDA-graph launch."""
        window = self.optimization_window()
        free = [k for k in window if k.index != 0]
        if not free:
            return None
        from . import ops
        from .primitives import pose_chain
        W, H = self.camera.width, self.camera.height
        gt = torch.stack([k.img for k in window])
        exposure = torch.stack([k.exposure for k in window])
        Ks = self.camera.intrinsics.expand(len(window), 3, 3).contiguous()
        P = [p.detach() for p in self.splats.params]
        bg = self.bg1.expand(len(window), 3).contiguous()
        nf = len(free)
        fixed = [k for k in window if k.index == 0]            # keyframe 0 sorts first in the window
        with torch.no_grad():
            Rt_all = torch.stack([k.pose.Rt for k in window])
            x0_dev = torch.cat([torch.cat([k.pose.dR, k.pose.dt]) for k in free])
            fixed_dR = torch.stack([k.pose.dR.detach() for k in fixed]) if fixed else None
            fixed_dt = torch.stack([k.pose.dt.detach() for k in fixed]) if fixed else None
        theta = torch.zeros(9 * nf, device=self.dev, requires_grad=True)
        result = torch.zeros(1 + 9 * nf, device=self.dev)
        theta_host = x0_dev.cpu().pin_memory()
        result_host = torch.zeros(1 + 9 * nf).pin_memory()
        theta_np, result_np = theta_host.numpy(), result_host.numpy()
        stream = torch.cuda.current_stream(self.dev)

        def fn():
            with torch.no_grad():
                theta.copy_(theta_host, non_blocking=True)
            theta.grad = None
            th = theta.view(nf, 9)
            dR, dt = th[:, :6], th[:, 6:]
            if fixed:
                dR, dt = torch.cat([fixed_dR, dR]), torch.cat([fixed_dt, dt])
            out = rasterization(P[0], P[1], P[2], P[3], P[4], pose_chain(Rt_all, dR, dt), Ks, W, H, packed=False,
                                render_mode="RGB+D", log_uncertainties=P[5], backgrounds=bg)
            total, _ = photometric_ssim_loss(out.render, gt, exposure, 0.0, "mapping")
            total.backward()
            torch.cat([total.detach().view(1), theta.grad], out=result)
            result_host.copy_(result, non_blocking=True)
            return total.detach()

        graph: List[Optional[GraphedStep]] = []

        def evaluate(x):
            theta_np[:] = x
            if not graph:
                # the first evaluation runs eagerly and sizes the key buffers of the capture; later
                # ones are one graph launch (upload, render + loss + backward to the poses, download)
                fn()
                stream.synchronize()
                first = float(result_np[0]), result_np[1:].copy()
                graph.append(GraphedStep(fn, headroom=1.5, n_isects_hint=int(ops.last_n_isects.item()),
                                         pool=shared_pool(self.dev)))
                self.stats.pose_lbfgs_evals += 1
                return first
            if graph[0] is not None:
                graph[0].replay()
                stream.synchronize()                       # the evaluation's sync
                if int(graph[0].overflow.item()) == 0:
                    self.stats.pose_lbfgs_evals += 1
                    return float(result_np[0]), result_np[1:]
                # a trial pose that sees more intersections than the captured buffers hold: redo it
                # with exact buffers and stay eager for the rest of this round
                graph[0] = None
                self.stats.pose_lbfgs_overflows += 1
            fn()
            stream.synchronize()
            self.stats.pose_lbfgs_evals += 1
            return float(result_np[0]), result_np[1:]

        res = lbfgs_minimize(evaluate, theta_np.copy(), lr=1.0, max_iter=20, tolerance_change=1e-7, history_size=10)
        with torch.no_grad():
            th = torch.from_numpy(np.asarray(res.x, dtype=np.float32)).to(self.dev)
            for i, k in enumerate(free):
                k.pose.dR.copy_(th[9 * i:9 * i + 6])
                k.pose.dt.copy_(th[9 * i + 6:9 * i + 9])
        return res.last_loss

    def sync(self) -> None:
        """Backend.sync -> Frontend.sync (backend.py:508-519, frontend.py:254-269)."""
        P = [p.detach() for p in self.splats.params]
        ages = torch.zeros(self.splats.n, dtype=torch.long, device=self.dev)
        self.tracker.set_map(GaussianSplattingData(P[0], P[1], P[2], P[3], P[4], P[5], ages))

    # ---- one frame --------------------------------------------------------------------------------
    def process_frame(self, img: Tensor, idle_rounds: int = 1, gt_depth: Optional[Tensor] = None) -> Tensor:
        """Track `img` [H,W,3] (device, values in [0,1]) and let the backend work; returns the frame's
        world->camera matrix."""
        from . import ops
        with ops.deterministic_mode(self.conf.deterministic):
            return self._process_frame(img, idle_rounds, gt_depth)

    def _process_frame(self, img: Tensor, idle_rounds: int, gt_depth: Optional[Tensor]) -> Tensor:
        idx = self.stats.frames
        sync = torch.cuda.synchronize
        if idx == 0:
            t0 = time.perf_counter()
            self.initialize(img, gt_depth)
            self.pause_map_optim = False
            self.optimize_map(self.conf.num_iters_initialization, prune=False, regularize=True)
            self.sync()
            sync()
            self.stats.t_mapping += time.perf_counter() - t0
            viewmat = torch.eye(4, device=self.dev)
            self.last_exposure = torch.zeros(2, device=self.dev)
        else:
            t0 = time.perf_counter()
            prev2 = self.frame_poses[-2] if len(self.frame_poses) > 1 else None
            init = constant_motion(prev2, self.frame_poses[-1])
            r = self.tracker.track_lbfgs(img, init, self.last_exposure,
                                         gt_depth if self.conf.use_gt_depths else None)
            viewmat, self.last_exposure = r.viewmat, r.exposure
            self.stats.tracking_evals += r.n_evals
            self.stats.tracking_syncs += r.n_syncs
            self.stats.last_tracking_loss = r.loss
            sync()
            t1 = time.perf_counter()
            self.stats.t_tracking += t1 - t0
            if self.to_insert_keyframe(viewmat):
                self.pause_map_optim = False
                self.add_keyframe(idx, img, viewmat, r.exposure, gt_depth)
                self.optimize_map(1, prune=True, regularize=False)
            for _ in range(idle_rounds):                  # Backend.run with an empty queue (:846-852)
                if self.pause_map_optim:
                    break
                self.optimize_map()
                if len(self.keyframes) > 1:
                    self.run_pruning()
                    if self.conf.optimize_poses:
                        self.optimize_poses_lbfgs()
            if idx % self.conf.sync_every == 0:
                self.sync()
            sync()
            self.stats.t_mapping += time.perf_counter() - t1
        self.frame_poses.append(viewmat)
        self.stats.frames += 1
        self.stats.n_splats = self.splats.n
        return viewmat


@torch.no_grad()
def synthetic_sequence(n_frames: int, n_gauss: int = 20000, scale: float = 1.0, device="cuda", seed: int = 0):
    """(camera, images [F,H,W,3], gt viewmats [F,4,4], depths [F,H,W]): renders of a random `synthetic.make_scene` map
    along `synthetic.smooth_trajectory`, produced by this package's own rasterizer (the data of cfg5)."""
    from .synthetic import make_scene, smooth_trajectory
    sc = make_scene(n_gauss, 1, "tum", scale, seed=seed).to(device)
    cam = Camera(sc.Ks[0].clone(), sc.height, sc.width)
    poses = smooth_trajectory(n_frames).to(device)
    imgs, depths = [], []
    P = sc.params()
    for i in range(0, n_frames, 8):
        vm = poses[i:i + 8].contiguous()
        C = vm.shape[0]
        out = rasterization(P[0], P[1], P[2], P[3], P[4], vm, sc.Ks[0].expand(C, 3, 3).contiguous(), sc.width,
                            sc.height, packed=False, render_mode="RGB+D", log_uncertainties=P[5],
                            backgrounds=torch.full((C, 3), 0.5, device=device))
        imgs.append(out.rgbs.clamp(0, 1).clone())
        depths.append(out.depthmaps.clone())          # accumulated (alpha-unnormalised) depth, like the map's
    return cam, torch.cat(imgs), poses, torch.cat(depths)


def run_sequence(n_frames: int, n_gauss: int = 20000, scale: float = 1.0, device="cuda", seed: int = 0,
                 map_conf: Optional[MapConfig] = None, track_conf: Optional[TrackingConfig] = None,
                 idle_rounds: int = 1, group=LOCAL):
    """cfg5: the whole loop over a synthetic sequence -> (Slam, frames/s, gt viewmats).  `group`: see Slam."""
    cam, imgs, gt_poses, depths = synthetic_sequence(n_frames, n_gauss, scale, device, seed)
    slam = Slam(cam, map_conf, track_conf, device, seed, group=group)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(n_frames):
        slam.process_frame(imgs[i], idle_rounds, depths[i])
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    return slam, n_frames / dt, gt_poses
