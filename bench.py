#!/usr/bin/env python
"""bench.py -- fwd+bwd rasterization throughput of the gslam hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Workload (BASELINE.json configs[1] shape): 200k Gaussians, 640x480 TUM-fr1
intrinsics, RGB + depth + beta channels, ONE view per GPU per step.  A step is
rasterization() forward, the active-NeRF photometric loss, and the backward
pass to every Gaussian parameter and to the camera viewmat.  Tracking has one
view and does not shard, so with N > 1 the ranks are independent replicas, each
on its own view (weak scaling, no data-path collective).  The `mapping` object
of the same line is the second half of the metric: the 8-keyframe mapping step
with the window's views sharded over the ranks and ONE NCCL all-reduce of the
flat Gaussian-parameter gradient per step (strong scaling).

Prints ONE JSON line (rank 0).  `--impl reference` times the CPU oracle
(oracle/torch_ref.py, the restatement of the reference's math: the reference's
own kernels live in un-vendored CUDA extensions and cannot run on a CPU) on the
box's host cores on a bounded sample of the same workload.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_GAUSS = 200_000
METRIC = "fwd+bwd rasterize it/s (640x480, 200k Gaussians)"
UNIT = "it/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--n-gauss", type=int, default=N_GAUSS)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--map-gauss", type=int, default=500_000)
    ap.add_argument("--no-mapping", action="store_true")
    ap.add_argument("--map-eager", action="store_true")
    ap.add_argument("--map-comm", default="auto", choices=["auto", "nccl", "nvls"])
    ap.add_argument("--map-big", default="auto", choices=["auto", "on", "off"],
                    help="the cfg4 mapping leg (3M Gaussians, SH3, 1280x720, 16 views); auto: at 8 GPUs")
    ap.add_argument("--big-gauss", type=int, default=3_000_000)
    ap.add_argument("--track-frames", type=int, default=12)
    ap.add_argument("--slam-frames", type=int, default=300)
    ap.add_argument("--no-sequence", action="store_true")
    ap.add_argument("--monocular", action="store_true",
                    help="also run the sequence leg without depth input (gauge-free: its error against the ground "
                         "truth is not meaningful on the synthetic scene)")
    ap.add_argument("--det", action="store_true", help="run every leg in deterministic accumulation mode")
    ap.add_argument("--no-f64", action="store_true", help="skip the float64 oracle run of the parity report")
    ap.add_argument("--leg-deadline", type=float, default=420.0,
                    help="seconds the secondary legs (mapping, sequences, CPU baseline) may take together before "
                         "rank 0 prints the headline line without them and every rank exits")
    ap.add_argument("--no-cfg1", action="store_true", help="skip the 10k-Gaussian 320x240 leg (BASELINE configs[0])")
    return ap.parse_args()


# the sources that define the kernels of the tracking step (what profiles/traffic.json quotes DRAM bytes for)
TRAFFIC_SOURCES = ("common.cuh", "composite.cuh", "composite.cu", "composite_bwd.cu", "project.cu", "binning.cu",
                   "loss.cu", "sh.cuh", "trace.cuh")


def csrc_sha() -> str:
    """Hash of the sources of the measured kernels: what a committed profile belongs to."""
    import hashlib
    h = hashlib.sha256()
    d = os.path.join(ROOT, "gslam_b200", "csrc")
    for fn in TRAFFIC_SOURCES:
        with open(os.path.join(d, fn), "rb") as f:
            h.update(fn.encode() + b"\0" + f.read())
    return h.hexdigest()[:12]


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured", float(d.get("sm_max_mhz", 1965.0))
    return 6650.0, "fallback", 1965.0


# ----------------------------------------------------------------------------------------------
# CPU arm: the oracle (oracle/torch_ref.py) on the SAME full view as the GPU arm, on the host cores
# ----------------------------------------------------------------------------------------------
def workload_string(n: int, world: int = 1, scale: float = 1.0) -> str:
    w, h = int(round(640 * scale)), int(round(480 * scale))
    return (f"cfg2-shape: 1 view per GPU, N={n}, {w}x{h} TUM-fr1, RGB+D+beta, "
            "fwd + active-NeRF loss + bwd (all Gaussian grads + viewmat grad)"
            + (", independent replicas (tracking does not shard)" if world > 1 else ""))


def cpu_oracle_step(sc, gt, dtype=None, stats: bool = False):
    """One fwd + active-NeRF tracking loss + bwd of the CPU oracle on the whole view ->
    (seconds, dict of outputs and gradients)."""
    import torch
    from oracle import torch_ref as R
    cast = (lambda t: t.to(dtype)) if dtype is not None else (lambda t: t)
    ps = [cast(p.clone()).requires_grad_(True) for p in sc.params()]
    vm = cast(sc.viewmats.clone()).requires_grad_(True)
    K, g = cast(sc.Ks), cast(gt)
    t0 = time.perf_counter()
    res = R.rasterization(ps[0], ps[1], ps[2], ps[3], ps[4], vm, K, sc.width, sc.height,
                          backgrounds=torch.zeros(1, 3, dtype=vm.dtype), render_mode="RGB+D",
                          log_uncertainties=ps[5], return_stats=stats)
    out, st = res if stats else (res, None)
    loss = R.tracking_loss(g, out.rgbs, out.betas)
    loss.backward()
    dt = time.perf_counter() - t0
    names = ["g_means", "g_quats", "g_log_scales", "g_logit_opacities", "g_logit_colors", "g_log_unc"]
    d = {"rgb": out.rgbs.detach(), "depth": out.depthmaps.detach(), "beta": out.betas.detach(),
         "alpha": out.alphas.detach(), "loss": loss.detach(), "g_viewmat": vm.grad[:, :3],
         "n_isects": int(out.flatten_ids.numel()), "pairs": st["P"] if st else None,
         "n_touched": out.n_touched, "radii": out.radii}
    d.update({nm: p.grad for nm, p in zip(names, ps)})
    return dt, d


def cpu_oracle_rate(n_gauss: int, reps: int = 1, warm: int = 0, scale: float = 1.0, keep: bool = False):
    """it/s of the CPU oracle on the FULL view (every step is the whole workload, nothing is scaled):
    -> (it/s, median seconds per step, cores, sample description, outputs of the last step or None)."""
    import torch
    from gslam_b200.synthetic import make_scene, target_images

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    sc = make_scene(n_gauss, 1, "tum", scale)
    gt = target_images(1, sc.height, sc.width)
    times, last = [], None
    for it in range(warm + reps):
        dt, d = cpu_oracle_step(sc, gt, stats=keep and it == warm + reps - 1)
        if it >= warm:
            times.append(dt)
        last = d if keep else None
    t = statistics.median(times)
    sample = (f"oracle/torch_ref.py fwd+loss+bwd on {n_gauss} Gaussians, the whole {sc.width}x{sc.height} view, "
              f"{reps} timed step(s) after {warm} warm-up, median; nothing scaled")
    return 1.0 / t, t, cores, sample, last


def run_reference(args):
    """The reference arm of this tier: the CPU oracle port on the host cores, the SAME workload as the
    GPU arm (whole 640x480 view, 200k Gaussians), exactly `warmup` untimed + `steps` timed steps."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    t0 = time.perf_counter()
    value, t, cores, sample, _ = cpu_oracle_rate(args.n_gauss, reps=args.steps, warm=args.warmup)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 / value,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": workload_string(args.n_gauss, 1),
                   "note": "CPU oracle (port of the reference math); reference kernels are un-vendored CUDA",
                   "extrapolated": False, "wall_s": None},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    line["config"]["wall_s"] = round(time.perf_counter() - t0, 1)
    print(json.dumps(line), flush=True)


def rel_err(a, b) -> float:
    """||a - b|| / ||b|| in float64 on the host."""
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).norm() / b.norm().clamp(min=1e-30))


PARITY_KEYS = ["rgb", "depth", "beta", "alpha", "g_means", "g_quats", "g_log_scales", "g_logit_opacities",
               "g_logit_colors", "g_log_unc", "g_viewmat"]
IMG_TOL, GRAD_TOL = 1e-4, 1e-3          # BASELINE.json north_star tolerances


def parity_report(gpu: dict, ref: dict, ref64: dict = None) -> dict:
    """Relative errors of the GPU step against the oracle's on the identical inputs; with the float64
    oracle also each side's error against it (so "GPU fp32 error <= oracle fp32 error" can be read off)."""
    rep = {k: rel_err(gpu[k], ref[k]) for k in PARITY_KEYS}
    rep["loss"] = abs(float(gpu["loss"]) - float(ref["loss"])) / max(abs(float(ref["loss"])), 1e-30)
    rep["n_isects_equal"] = gpu["n_isects"] == ref["n_isects"]
    rep["pairs_equal"] = None if ref.get("pairs") is None or gpu.get("pairs") is None else gpu["pairs"] == ref["pairs"]
    rep["n_isects"] = [gpu["n_isects"], ref["n_isects"]]
    rep["pairs"] = [gpu.get("pairs"), ref.get("pairs")]
    nt_g, nt_r = gpu["n_touched"].cpu().long(), ref["n_touched"].long()
    rep["n_touched_mismatch_frac"] = float((nt_g != nt_r).float().mean())
    rep["radii_mismatch_frac"] = float((gpu["radii"].cpu() != ref["radii"]).float().mean())
    if ref64 is not None:
        rep["vs_f64"] = {k: {"gpu_f32": rel_err(gpu[k], ref64[k]), "oracle_f32": rel_err(ref[k], ref64[k])}
                         for k in PARITY_KEYS}
    worst_img = max(rep[k] for k in PARITY_KEYS[:4])
    worst_grad = max(rep[k] for k in PARITY_KEYS[4:])
    rep["tolerance"] = {"images": IMG_TOL, "gradients": GRAD_TOL}
    rep["ok"] = bool(worst_img <= IMG_TOL and worst_grad <= GRAD_TOL)
    return rep


# ----------------------------------------------------------------------------------------------
# GPU arm
# ----------------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.idx = gpu_index
        self.p = None

    def start(self):
        try:
            self.p = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200",
                 "-i", str(self.idx)], stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.p.terminate()
        self.p.wait()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.f.read().splitlines():
            c = [x.strip() for x in ln.split(",")]
            if len(c) < 9:
                continue
            try:
                sm.append(float(c[1])); mx.append(float(c[2]))
            except ValueError:
                continue
            for nm, v in zip(names, c[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        os.unlink(self.f.name)
        return {"sm_mhz": statistics.median(sm) if sm else None,
                "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons), "samples": len(sm)}


def mapping_leg(args, dev, rank, world):
    """Second half of BASELINE.json's metric: mapping views/s (configs[2] shape: 8-keyframe window,
    500k Gaussians, render + photometric/SSIM losses + backward + gradient all-reduce + Adam +
    opacity decay = the body of Backend.optimize_map, gslam/backend.py:260-359), the window's
    views sharded window[rank::world] (strong scaling: the window is fixed at 8 views)."""
    import torch
    import torch.distributed as dist
    from gslam_b200.mapping import DEFAULT_LRS, PARAM_NAMES, FlatSplats, MappingStep, shard_views
    from gslam_b200.optim import FusedAdam
    from gslam_b200.synthetic import make_scene, target_images

    n_views, n = 8, args.map_gauss
    sc = make_scene(n, n_views, "tum", 1.0)
    mine = shard_views(n_views, rank, world)
    if not mine:
        mine = [rank % n_views]          # more ranks than views: replicate (never at 1/2/4/8)
    g = sc.to(dev)
    want_nvls = world > 1 and args.map_comm != "nccl"
    splats = FlatSplats(g.params(), align=4 * world, symmetric=dist.group.WORLD if want_nvls else None)
    opt = FusedAdam([dict(params=[p], lr=DEFAULT_LRS[nm]) for p, nm in zip(splats.params, PARAM_NAMES)],
                    capturable=True)
    mstep = MappingStep(splats, opt, comm=args.map_comm)
    vm, K = g.viewmats[mine].contiguous(), g.Ks[mine].contiguous()
    gt = target_images(n_views, sc.height, sc.width)[mine].to(dev)
    exposure = torch.zeros(len(mine), 2, device=dev)
    steps = max(3, min(args.steps, 20))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def eager():
        return mstep(vm, K, gt, exposure, sc.width, sc.height, n_views).total

    for _ in range(3):
        eager()
    barrier()
    # phase times of the step's tail from CUDA events of eager steps (max over ranks): the reduce-scatter of
    # the gradient rows, the fused update of this rank's Gaussian range, the all-gather of the parameters
    phases = {"reduce_scatter_ms": [], "update_ms": [], "all_gather_ms": [], "render_loss_backward_ms": []}
    for _ in range(5):
        mstep.phase_events = []
        eager()
        torch.cuda.synchronize()
        ev_ = dict(mstep.phase_events)
        phases["render_loss_backward_ms"].append(ev_["begin"].elapsed_time(ev_["backward_done"]))
        phases["reduce_scatter_ms"].append(ev_["backward_done"].elapsed_time(ev_["reduce_scatter_done"]))
        phases["update_ms"].append(ev_["reduce_scatter_done"].elapsed_time(ev_["update_done"]))
        phases["all_gather_ms"].append(ev_["update_done"].elapsed_time(ev_["all_gather_done"]))
    mstep.phase_events = None
    ph = torch.tensor([statistics.median(v) for v in phases.values()], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ph, op=dist.ReduceOp.MAX)
    phase_ms = {k: round(float(x), 4) for k, x in zip(phases, ph)}
    barrier()
    # the step is launch-bound at one view per GPU (about 60 launches), so it is replayed as ONE CUDA
    # graph (NCCL all-reduces included); --map-eager times the plain launches instead
    run, mode = eager, "eager launches"
    if not args.map_eager:
        from gslam_b200.graph import GraphedStep
        g_step = GraphedStep(eager)
        run, mode = g_step.replay, "CUDA-graph replay"
        for _ in range(3):
            run()
        barrier()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    for e0, e1 in ev:
        e0.record()
        run()
        e1.record()
    barrier()
    if not args.map_eager:
        g_step.check()
    ms = sum(a.elapsed_time(b) for a, b in ev) / steps
    t = torch.tensor([ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    return {"metric": "mapping views/s (8-keyframe window, full optimize_map step)", "value": n_views * 1e3 / ms,
            "unit": "views/s", "ms_per_step": ms, "steps": steps, "n_views": n_views, "n_gauss": n,
            "views_per_rank": len(mine), "scaling": "strong", "n_ranks": world,
            "comm": "none (one rank)" if world == 1 else (
                "nvls: one kernel = in-switch reduce-scatter (multimem.ld_reduce) + update + multicast all-gather "
                "(multimem.st), two device barriers" if mstep._nvls is not None else
                "nccl: reduce_scatter_tensor + update kernel + coalesced all_gather_into_tensor"),
            "collective_ms": round(phase_ms["reduce_scatter_ms"] + phase_ms["all_gather_ms"], 4),
            "post_reduce_ms": phase_ms["update_ms"], "eager_phases": phase_ms,
            "step": "render RGB+D+beta, exposure + active-NeRF NLL + 0.2 SSIM(valid), backward to gradient rows, "
                    "reduce-scatter, fused (isotropic reg + Adam + opacity decay) on this rank's Gaussians, "
                    "all-gather of the parameters", "timed_loop": mode}


def mapping_big_leg(args, dev, rank, world):
    """BASELINE.json configs[3]: 3M Gaussians, SH degree 3 ([N,16,3] coefficients), 1280x720 OAK-D-shaped
    frames, 16-keyframe window sharded over the ranks (2 views per GPU on 8), the same step as mapping_leg.
    192 MB of parameters and 768 MB of gradient rows per rank: the first place the NVLink BANDWIDTH of the
    reduce-scatter / all-gather shows, not its latency."""
    import torch
    import torch.distributed as dist
    from gslam_b200.graph import GraphedStep
    from gslam_b200.mapping import DEFAULT_LRS, PARAM_NAMES, FlatSplats, MappingStep, shard_views
    from gslam_b200.optim import FusedAdam
    from gslam_b200.synthetic import make_scene, target_images

    n_views, n = 16, args.big_gauss
    sc = make_scene(n, n_views, "oakd", 1.0)
    mine = shard_views(n_views, rank, world)
    tensors = [t.to(dev) for t in sc.params()]
    gsh = torch.Generator().manual_seed(7)
    sh = 0.1 * torch.randn(n, 16, 3, generator=gsh)
    sh[:, 0] = sc.logit_colors * 0.5            # DC term from the logit colours
    tensors[4] = sh.to(dev)
    del sh
    want_nvls = world > 1 and args.map_comm != "nccl"
    splats = FlatSplats(tensors, align=4 * world, symmetric=dist.group.WORLD if want_nvls else None)
    del tensors
    opt = FusedAdam([dict(params=[p], lr=DEFAULT_LRS[nm]) for p, nm in zip(splats.params, PARAM_NAMES)],
                    capturable=True)
    mstep = MappingStep(splats, opt, comm=args.map_comm, sh_degree=3)
    vm, K = sc.viewmats[mine].contiguous().to(dev), sc.Ks[mine].contiguous().to(dev)
    gt = target_images(n_views, sc.height, sc.width)[mine].to(dev)
    exposure = torch.zeros(len(mine), 2, device=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def eager():
        return mstep(vm, K, gt, exposure, sc.width, sc.height, n_views).total

    for _ in range(2):
        eager()
    barrier()
    from gslam_b200 import ops
    n_isects = int(ops.last_n_isects.item())
    g_step = GraphedStep(eager)
    for _ in range(2):
        g_step.replay()
    barrier()
    steps = 5
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    for e0, e1 in ev:
        e0.record()
        g_step.replay()
        e1.record()
    barrier()
    g_step.check()
    ms = sum(a.elapsed_time(b) for a, b in ev) / steps
    t = torch.tensor([ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    res = {"metric": "mapping views/s (cfg4: 16-keyframe window, 3M Gaussians, SH3, 1280x720)",
           "value": n_views * 1e3 / ms, "unit": "views/s", "ms_per_step": ms, "steps": steps, "n_views": n_views,
           "n_gauss": n, "sh_degree": 3, "views_per_rank": len(mine), "n_ranks": world,
           "n_isects_rank0": n_isects, "row_floats": 64, "gradient_rows_mb_per_rank": round(n * 64 * 4 / 1e6, 1),
           "parameter_mb": round(n * 60 * 4 / 1e6, 1),
           "comm": "none (one rank)" if world == 1 else ("nvls fused kernel" if mstep._nvls is not None else "nccl"),
           "timed_loop": "CUDA-graph replay"}
    del g_step, mstep, opt, splats
    torch.cuda.empty_cache()
    return res


def tracking_closure_leg(args, dev):
    """The reference's own tracking iteration (gslam/frontend.py:604-662, the Adam phase of
    igs_track_lbfgs): the map is frozen (no_grad clone, gslam/backend.py:515), only the frame's pose
    (PoseZhou dR, dt) and exposure receive gradients; closure = splats([camera], [pose],
    render_depth=True) -> exposure + active-NeRF tracking loss -> backward -> Adam step."""
    import torch
    from gslam_b200.graph import GraphedStep
    from gslam_b200.losses import tracking_loss
    from gslam_b200.map import GaussianSplattingData
    from gslam_b200.optim import FusedAdam
    from gslam_b200.primitives import Camera, PoseZhou
    from gslam_b200.synthetic import make_scene, target_images

    n = args.n_gauss
    sc = make_scene(n, 1, "tum", 1.0).to(dev)
    p = sc.params()
    splats = GaussianSplattingData(p[0], p[1], p[2], p[3], p[4], p[5],
                                   torch.zeros(n, dtype=torch.long, device=dev)).no_grad_clone()
    cam = Camera(sc.Ks[0].clone(), sc.height, sc.width)
    pose = PoseZhou(sc.viewmats[0].clone()).to(dev)
    exposure = torch.zeros(2, device=dev, requires_grad=True)
    gt = target_images(1, sc.height, sc.width).to(dev)
    opt = FusedAdam([pose.dR, pose.dt, exposure], lr=3e-3, capturable=True)

    def step():
        for q in (pose.dR, pose.dt, exposure):
            q.grad = None
        out = splats([cam], [pose], render_depth=True)
        loss = tracking_loss(out.render, gt, exposure, "active-nerf", unit_upstream=True)
        loss.backward()
        opt.step()
        return loss

    gs = GraphedStep(step)
    for _ in range(3):
        gs.replay()
    torch.cuda.synchronize()
    steps = args.steps
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    for e0, e1 in ev:
        flush.zero_()
        e0.record()
        gs.replay()
        e1.record()
    torch.cuda.synchronize()
    gs.check()
    ms = sum(a.elapsed_time(b) for a, b in ev) / steps
    return {"metric": "tracking closure it/s (frozen map, pose + exposure gradients, Adam step)",
            "value": 1e3 / ms, "unit": "it/s", "ms_per_step": ms, "steps": steps, "n_gauss": n,
            "step": "PoseZhou chain, render RGB+D+beta, exposure + active-NeRF loss, backward to dR/dt/exposure, "
                    "fused Adam", "timed_loop": "CUDA-graph replay, L2 flushed between steps",
            "launch_calls_per_step": gs.launches_per_replay}


def tracking_frames_leg(args, dev):
    """BASELINE cfg2 as the reference really schedules it (gslam/frontend.py:604-662): per frame 10
    Adam steps + torch L-BFGS (strong Wolfe, <= 25 evaluations) on a frozen 200k-Gaussian map,
    gslam_b200.tracking.Tracker; wall clock per frame (the L-BFGS phase is host-driven), every
    frame a new image and the constant-motion start."""
    import time
    import torch
    from gslam_b200.map import GaussianSplattingData
    from gslam_b200.primitives import Camera
    from gslam_b200.synthetic import make_scene, smooth_trajectory
    from gslam_b200.tracking import Tracker, TrackingConfig, constant_motion
    from gslam_b200 import _lib

    n, F = args.n_gauss, args.track_frames
    sc = make_scene(n, 1, "tum", 1.0).to(dev)
    splats = GaussianSplattingData(*sc.params(), torch.zeros(n, dtype=torch.long, device=dev)).no_grad_clone()
    cam = Camera(sc.Ks[0].clone(), sc.height, sc.width)
    poses = smooth_trajectory(F + 2).to(dev)
    with torch.no_grad():
        imgs = [splats.render([cam], [poses[i]]).rgbs[0].clamp(0, 1).clone() for i in range(F + 2)]
    # the frames arrive from the host: pinned buffers, uploaded inside the timed loop; the pose goes back
    host_imgs = [im.cpu().pin_memory() for im in imgs]
    img_dev = torch.empty_like(imgs[0])
    pose_host = torch.empty(4, 4).pin_memory()
    tr = Tracker(cam, TrackingConfig(), dev)
    tr.set_map(splats)
    est = [poses[0].clone()]
    est.append(tr.track_lbfgs(imgs[1], est[0]).viewmat)      # untimed: captures both graphs
    est.append(tr.track_lbfgs(imgs[1], constant_motion(est[0], est[1])).viewmat)   # untimed warm-up frame
    est = est[:2]
    torch.cuda.synchronize()
    evals = syncs = 0
    launches0 = _lib.launch_count
    t0 = time.perf_counter()
    for i in range(2, F + 2):
        img_dev.copy_(host_imgs[i], non_blocking=True)
        r = tr.track_lbfgs(img_dev, constant_motion(est[-2], est[-1]))
        pose_host.copy_(r.viewmat, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        est.append(r.viewmat)
        evals += r.n_evals
        syncs += r.n_syncs + 1
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    err = float((torch.stack(est[2:])[:, :3, 3] - poses[2:, :3, 3]).norm(dim=1).mean())
    return {"metric": "tracking frames/s (10 Adam + L-BFGS strong-Wolfe per frame, frozen map)",
            "value": F / dt, "unit": "frames/s", "frames": F, "n_gauss": n, "evals_per_frame": evals / F,
            "host_syncs_per_frame": syncs / F, "ms_per_eval": 1e3 * dt / max(evals, 1),
            "mean_translation_error_m": err, "h2d_bytes_per_frame": host_imgs[0].numel() * 4,
            "d2h_bytes_per_frame": 64 + 48 * (evals - 10 * F) // F,
            "timed_loop": "wall clock, host image in (pinned) -> pose out; graph replay per evaluation"}


def slam_sequence_leg(args, dev, rgbd: bool, group="local"):
    """BASELINE cfg5: synthetic sequence through tracking + mapping + insertion / pruning
    (gslam_b200.slam, the inline equivalent of main.py), frames/s by wall clock.  rgbd=True is the
    reference's use_gt_depths mode: its poses are comparable with the ground truth (gauge fixed by
    frame 0), so the trajectory error is reported; the monocular default has the usual scale freedom."""
    import torch
    from gslam_b200.slam import MapConfig, run_sequence
    F = args.slam_frames
    # deterministic accumulation: the same sequence gives the same keyframes, map and trajectory in every run
    slam, fps, gt = run_sequence(F, n_gauss=20_000, scale=1.0, device=dev,
                                 map_conf=MapConfig(use_gt_depths=rgbd, deterministic=True), group=group)
    st = slam.stats
    est = torch.stack(slam.frame_poses)
    err = (est[:, :3, 3] - gt[:, :3, 3]).norm(dim=1)
    return {"metric": "sequence frames/s (tracking + mapping + insertion/pruning, 640x480)", "value": fps,
            "unit": "frames/s", "mode": "rgbd (use_gt_depths)" if rgbd else "monocular", "frames": F,
            "deterministic": True, "mapping_ranks": slam.world,
            "keyframes": st.keyframes, "n_splats": st.n_splats,
            "mapping_iters": st.mapping_iters, "tracking_evals": st.tracking_evals,
            "pose_lbfgs_evals": st.pose_lbfgs_evals, "inserted": st.inserted, "pruned": st.pruned,
            "overflow_rollbacks": st.mapping_overflow_rollbacks,
            "s_tracking": st.t_tracking, "s_mapping": st.t_mapping,
            "translation_error_m": {"mean": float(err.mean()), "max": float(err.max()),
                                    "path_length": float((gt[1:, :3, 3] - gt[:-1, :3, 3]).norm(dim=1).sum())},
            "lost_track": bool(float(err.mean()) > 0.25 * float((gt[1:, :3, 3] - gt[:-1, :3, 3]).norm(dim=1).sum())),
            "includes": "frame 0 initialisation (up to 400 mapping iterations)"}


def cfg1_leg(dev):
    """BASELINE.json configs[0]: the CPU oracle's own case -- 10k Gaussians, 320x240, one view, fwd + loss +
    bwd -- timed on the host cores (median of 5 after 1 warm-up) with the GPU path beside it at the same
    shape (CUDA-graph replay, L2 flushed between steps), and the parity of the two."""
    import torch
    from gslam_b200.graph import GraphedStep
    from gslam_b200.losses import tracking_loss
    from gslam_b200.rasterization import rasterization
    from gslam_b200.synthetic import make_scene, target_images
    n, scale = 10_000, 0.5
    v, tsec, cores, sample, ref = cpu_oracle_rate(n, reps=5, warm=1, scale=scale, keep=True)
    sc = make_scene(n, 1, "tum", scale)
    g = sc.to(dev)
    params = [p.clone().requires_grad_(True) for p in g.params()]
    vm = g.viewmats.clone().requires_grad_(True)
    gt = target_images(1, sc.height, sc.width).to(dev)
    bg = torch.zeros(1, 3, device=dev)
    box = {}

    def step():
        for p in params:
            p.grad = None
        vm.grad = None
        out = rasterization(params[0], params[1], params[2], params[3], params[4], vm, g.Ks, sc.width, sc.height,
                            packed=False, render_mode="RGB+D", log_uncertainties=params[5], backgrounds=bg)
        loss = tracking_loss(out.render, gt, None, "active-nerf", unit_upstream=True)
        loss.backward()
        if box.get("keep"):
            box["out"] = out
        return loss

    box["keep"] = True
    loss = step()
    out = box.pop("out")
    box["keep"] = False
    names = ["g_means", "g_quats", "g_log_scales", "g_logit_opacities", "g_logit_colors", "g_log_unc"]
    gpu = {"rgb": out.rgbs.detach().cpu(), "depth": out.depthmaps.detach().cpu(), "beta": out.betas.detach().cpu(),
           "alpha": out.alphas.detach().cpu(), "loss": loss.detach().cpu(), "g_viewmat": vm.grad[:, :3].cpu(),
           "n_isects": int(out.flatten_ids.numel()), "pairs": None, "n_touched": out.n_touched.cpu(),
           "radii": out.radii.cpu()}
    gpu.update({nm: p.grad.cpu() for nm, p in zip(names, params)})
    # nothing of the eager step may stay alive into the capture: a live output keeps the parameters'
    # AccumulateGrad nodes, and the stream they ran on, in the graph (cudaErrorStreamCaptureImplicit)
    del out, loss
    gs = GraphedStep(step)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    steps = 30
    for _ in range(5):
        gs.replay()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    for e0, e1 in ev:
        flush.zero_()
        e0.record()
        gs.replay()
        e1.record()
    torch.cuda.synchronize()
    gs.check()
    ms = sum(a.elapsed_time(b) for a, b in ev) / steps
    return {"workload": workload_string(n, 1, scale) + " (BASELINE.json configs[0])",
            "cpu_oracle": {"value": v, "unit": UNIT, "cores": cores, "seconds": tsec, "sample": sample},
            "gpu": {"value": 1e3 / ms, "unit": UNIT, "ms_per_step": ms, "steps": steps,
                    "timed_loop": "CUDA-graph replay, L2 flushed between steps"},
            "parity": parity_report(gpu, ref)}


_T0 = time.perf_counter()


def progress(msg: str) -> None:
    """One line per leg on stderr (rank 0): where a run spends its wall time, and where it stopped if it did."""
    if int(os.environ.get("RANK", "0")) == 0:
        print(f"[bench {time.perf_counter() - _T0:7.1f} s] {msg}", file=sys.stderr, flush=True)


def run_ours(args):
    import torch
    import torch.distributed as dist
    from gslam_b200 import _lib, ops
    from gslam_b200.rasterization import rasterization
    from gslam_b200.synthetic import make_scene, target_images  # noqa: F811

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback in the product path)"
    ops.deterministic = bool(args.det)
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    n = args.n_gauss
    sc = make_scene(n, world, "tum", 1.0)       # one view of the mapping window per rank
    H, W = sc.height, sc.width
    # The step's inputs live in ONE pinned host arena and ONE flat device buffer (every tensor a 64-byte aligned
    # view): the end-to-end loop then uploads a step's inputs with one 15.7 MB copy instead of nine.
    srcs_cpu = [p.clone() for p in sc.params()] + [sc.viewmats[rank:rank + 1].clone(), sc.Ks[rank:rank + 1].clone(),
                                                   target_images(world, H, W)[rank:rank + 1].clone()]
    offs, total = [], 0
    for t_ in srcs_cpu:
        offs.append(total)
        total += -(-t_.numel() // 16) * 16
    host_arena = torch.empty(total, dtype=torch.float32).pin_memory()
    dev_arena = torch.empty(total, dtype=torch.float32, device=dev)

    def views(arena):
        return [arena[o:o + t_.numel()].view(t_.shape) for o, t_ in zip(offs, srcs_cpu)]
    host_views = views(host_arena)
    for hv, t_ in zip(host_views, srcs_cpu):
        hv.copy_(t_)
    dev_arena.copy_(host_arena)
    host, (host_vm, host_K, host_gt) = host_views[:6], host_views[6:]
    dev_views = views(dev_arena)
    params = [v.requires_grad_(True) for v in dev_views[:6]]
    vm = dev_views[6].requires_grad_(True)
    K, gt = dev_views[7], dev_views[8]
    bg = torch.zeros(1, 3, device=dev)
    flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device=dev)   # 2x the 126 MB L2

    from gslam_b200.losses import tracking_loss

    def loss_fn(out, gt_):
        # active-NeRF photometric term of the tracking step (gslam/frontend.py:113-129), fused kernel
        return tracking_loss(out.render, gt_, None, "active-nerf", unit_upstream=True)   # backward() starts here

    def step(ps, vm_, K_, gt_):
        for p in ps:
            p.grad = None
        vm_.grad = None
        out = rasterization(ps[0], ps[1], ps[2], ps[3], ps[4], vm_, K_, W, H, packed=False,
                            render_mode="RGB+D", log_uncertainties=ps[5], backgrounds=bg)
        loss = loss_fn(out, gt_)
        loss.backward()
        # tracking (one view) does not shard: with N > 1 the ranks are independent replicas, each on
        # its own view, and there is no data-path collective (the mapping leg has the real one)
        return loss, out

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step(params, vm, K, gt)
    barrier()

    progress('warm-up done, capturing the step')
    # The step is launch-bound in eager mode (about 50 launches + one host sync), so the timed
    # loop replays a CUDA graph of the SAME step() captured through the public API.
    from gslam_b200.graph import GraphedStep
    gstep = GraphedStep(lambda: step(params, vm, K, gt)[0])
    static_vm_grad = vm.grad          # graph-pool tensor refilled by every replay
    for _ in range(3):
        gstep.replay()
    barrier()

    # ---- timed region: K steps, one CUDA-event pair per step, L2 flushed between steps ----------
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
          for _ in range(args.steps)]
    barrier()
    wall0 = time.perf_counter()
    for e0, e1 in ev:
        flush_buf.zero_()
        e0.record()
        gstep.replay()
        e1.record()
    barrier()
    wall = time.perf_counter() - wall0
    gstep.check()
    launches = gstep.launches_per_replay * args.steps
    clocks = sampler.stop() if rank == 0 else None
    ms = sum(a.elapsed_time(b) for a, b in ev) / args.steps
    t = torch.tensor([ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    value = world * 1e3 / ms

    # eager (no graph) rate of the same step, for reference
    ev3 = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
           for _ in range(args.steps)]
    for e0, e1 in ev3:
        flush_buf.zero_()
        e0.record()
        step(params, vm, K, gt)
        e1.record()
    barrier()
    eager_ms = sum(a.elapsed_time(b) for a, b in ev3) / args.steps

    # ---- per-kernel durations (CUDA events around each C-ABI call, same steps, live) -------------
    ops.pair_counter = torch.zeros(1, dtype=torch.int64, device=dev)
    loss_g, out = step(params, vm, K, gt)
    torch.cuda.synchronize()
    P = int(ops.pair_counter.item())
    I = int(out.isect_ids.numel())
    ops.pair_counter = None
    # what the parity report compares with the oracle's step on the identical inputs (cpu_baseline leg)
    gpu_step = None
    if world == 1 and not args.no_cpu_baseline:
        gnames = ["g_means", "g_quats", "g_log_scales", "g_logit_opacities", "g_logit_colors", "g_log_unc"]
        gpu_step = {"rgb": out.rgbs.detach().cpu(), "depth": out.depthmaps.detach().cpu(),
                    "beta": out.betas.detach().cpu(), "alpha": out.alphas.detach().cpu(),
                    "loss": loss_g.detach().cpu(), "g_viewmat": vm.grad[:, :3].cpu(), "n_isects": I, "pairs": P,
                    "n_touched": out.n_touched.cpu(), "radii": out.radii.cpu()}
        gpu_step.update({nm: p.grad.cpu() for nm, p in zip(gnames, params)})
    del out
    # per-kernel pass: an event pair around every C-ABI call of live steps.  The host must run AHEAD of
    # the device for such a pair to bracket the kernel alone -- otherwise it also spans the launch
    # latency of a kernel that reaches an idle GPU (+5 us on every small kernel).  So the key buffers
    # get a fixed capacity (no mid-step host sync) and each step is queued behind a ~1.5 ms spin kernel.
    ops.isect_capacity = int(I * 1.3) + 4096
    try:
        _lib.start_timing()
        for _ in range(min(args.steps, 20)):
            flush_buf.zero_()
            torch.cuda._sleep(3_000_000)
            step(params, vm, K, gt)
        kt = _lib.stop_timing()
    finally:
        ops.isect_capacity = None

    progress('timed loop and per-kernel pass done')
    # ---- e2e: host buffers in, loss + pose gradient out, copies inside the timed region ----------
    h2d = host_arena.numel() * 4        # the pipelined loop copies the whole arena (64-byte padding included)
    res_host = torch.empty(17, dtype=torch.float32).pin_memory()
    d2h = res_host.numel() * 4

    def e2e_step():
        with torch.no_grad():
            for p, h in zip(params, host):
                p.copy_(h, non_blocking=True)
            vm.copy_(host_vm, non_blocking=True)
            K.copy_(host_K, non_blocking=True)
            gt.copy_(host_gt, non_blocking=True)
        loss = gstep.replay()
        res = torch.cat([loss.detach().view(1), static_vm_grad.view(-1)])
        res_host.copy_(res, non_blocking=True)
        torch.cuda.current_stream().synchronize()

    for _ in range(3):
        e2e_step()
    barrier()
    ev2 = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
           for _ in range(args.steps)]
    for e0, e1 in ev2:
        flush_buf.zero_()
        e0.record()
        e2e_step()
        e1.record()
    barrier()
    ms2_serial = sum(a.elapsed_time(b) for a, b in ev2) / args.steps

    # The same step as a software pipeline.  Step i+1's inputs travel host -> staging on a copy stream (ONE copy of
    # the pinned arena) while step i computes; the compute stream moves staging -> the step's tensors (one
    # device-to-device copy), replays the graph and copies the result to pinned memory; the host reads step i's
    # result after it has queued step i+1 (two result buffers), so the GPU never waits for the host's wake-up.
    # Every step still uploads all of its inputs and returns its result inside its own event pair: the upload of
    # step i+1 may only start after step i's e0 (it waits for the L2 flush before it) and step i's e1 waits for it,
    # so no transfer hides in the un-timed flush between the pairs.
    copy_stream = torch.cuda.Stream(device=dev)
    staging = [torch.empty_like(dev_arena) for _ in range(2)]
    res_hosts = [torch.empty(17, dtype=torch.float32).pin_memory() for _ in range(2)]
    up_done = [torch.cuda.Event() for _ in range(2)]
    slot_free = [torch.cuda.Event() for _ in range(2)]
    res_done = [torch.cuda.Event() for _ in range(2)]
    gate = torch.cuda.Event()
    consumed = []

    def upload(slot):
        with torch.cuda.stream(copy_stream), torch.no_grad():
            copy_stream.wait_event(gate)
            copy_stream.wait_event(slot_free[slot])
            staging[slot].copy_(host_arena, non_blocking=True)
            up_done[slot].record(copy_stream)

    def e2e_pipelined(i):
        slot = i & 1
        main = torch.cuda.current_stream()
        gate.record(main)                      # after this step's flush and e0
        upload(slot ^ 1)                       # next step's inputs
        with torch.no_grad():
            dev_arena.copy_(staging[slot], non_blocking=True)     # its upload was awaited at the end of step i-1
        slot_free[slot].record(main)
        loss = gstep.replay()
        res = torch.cat([loss.detach().view(1), static_vm_grad.view(-1)])
        res_hosts[slot].copy_(res, non_blocking=True)
        res_done[slot].record(main)
        main.wait_event(up_done[slot ^ 1])     # the next step's upload belongs to this step's timed region

    def consume(i):                            # the host's read of step i's result
        res_done[i & 1].synchronize()
        consumed.append(float(res_hosts[i & 1][0]))

    main0 = torch.cuda.current_stream()
    for sl in range(2):
        slot_free[sl].record(main0)
    gate.record(main0)
    upload(0)
    main0.wait_event(up_done[0])
    for i in range(4):
        e2e_pipelined(i)
        if i:
            consume(i - 1)
    consume(3)
    barrier()
    ev2 = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
           for _ in range(args.steps)]
    for i, (e0, e1) in enumerate(ev2):
        flush_buf.zero_()
        e0.record()
        e2e_pipelined(i)
        e1.record()
        if i:
            consume(i - 1)
    consume(args.steps - 1)
    barrier()
    copy_stream.synchronize()
    assert all(v == v for v in consumed)
    ea, eb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ea.record()
    for _ in range(5):
        staging[0].copy_(host_arena, non_blocking=True)
    eb.record()
    torch.cuda.synchronize()
    h2d_ms = ea.elapsed_time(eb) / 5           # the PCIe floor of a step: one arena copy on an idle link
    ms2 = sum(a.elapsed_time(b) for a, b in ev2) / args.steps
    t = torch.tensor([ms2, ms2_serial], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = world * 1e3 / float(t[0].item())
    e2e_serial = world * 1e3 / float(t[1].item())

    progress('e2e loops done')
    # The legs below (mapping, sequences) are secondary: if one of them wedges -- a collective that a rank never
    # joins -- the headline must still come out.  Every rank arms the same deadline; when it fires rank 0 prints the
    # line with what is measured so far and every rank leaves.
    import threading

    def bail():
        if rank == 0:
            fallback = {
                "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": workload_string(n, world),
                           "l2": "256 MiB memset between timed steps, outside each step's event pair"},
                "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                        "serial_value": e2e_serial},
                "gpu_launches": launches, "clocks": clocks,
                "error": f"a secondary leg did not finish within {args.leg_deadline} s (see the [bench] lines on "
                         "stderr for the last one that did); headline, e2e and the legs listed here are complete"}
            full = partial.pop("line", None)
            if full is not None:                # everything but the last leg (mapping_big) is in: print the whole line
                fallback = dict(full, error=fallback["error"])
                fallback["mapping_big"] = {"error": "did not finish before the deadline"}
            fallback.update(partial)            # the legs that did finish (the mapping object, if there, goes last)
            if "mapping" in fallback:
                fallback["mapping"] = fallback.pop("mapping")
            print(json.dumps(fallback), flush=True)
        sys.stderr.flush()
        os._exit(0)
    partial = {}
    watchdog = threading.Timer(args.leg_deadline, bail)
    watchdog.daemon = True
    watchdog.start()
    mapping = None
    if not args.no_mapping:
        mapping = mapping_leg(args, dev, rank, world)
    partial["mapping"] = mapping
    progress('mapping leg done')
    closure = tracking_closure_leg(args, dev) if rank == 0 else None
    import traceback
    seq_sharded = None
    if world > 1 and not args.no_sequence:
        # cfg5 with the mapping window sharded by view over ALL ranks (every rank runs the loop)
        try:
            seq_sharded = slam_sequence_leg(args, dev, True, group=None)
        except Exception as e:   # noqa: BLE001
            seq_sharded = {"error": f"{type(e).__name__}: {e}", "where": traceback.format_exc().splitlines()[-12:]}
    partial["sequence_sharded"] = seq_sharded
    progress('sharded sequence leg done' if seq_sharded is not None else 'sharded sequence leg skipped')
    track_frames = slam_seq = None
    if rank == 0 and not args.no_sequence:
        # secondary legs: a failure here is reported in place and must not take the headline line down
        try:
            track_frames = tracking_frames_leg(args, dev)
        except Exception as e:   # noqa: BLE001
            track_frames = {"error": f"{type(e).__name__}: {e}", "where": traceback.format_exc().splitlines()[-12:]}
        progress('tracking_frames leg done')
        slam_seq = {}
        for mode, rgbd in (("rgbd", True), ("monocular", False)) if args.monocular else (("rgbd", True),):
            try:
                slam_seq[mode] = slam_sequence_leg(args, dev, rgbd)
            except Exception as e:   # noqa: BLE001
                slam_seq[mode] = {"error": f"{type(e).__name__}: {e}", "where": traceback.format_exc().splitlines()[-12:]}

    partial.update({"tracking_closure": closure, "tracking_frames": track_frames, "sequence": slam_seq})
    progress('sequence legs done')
    mapping_big = None
    if rank == 0:
        hbm, which, sm_max = peaks()
        CN, px = n, H * W
        # algorithmic bytes per launch, C = 1 view per rank
        alg = {
            "gsb_project_fwd": 60 * n + 64 * CN,
            "gsb_project_bwd": 60 * n + (48 + 8 + 4 + 48) * CN + 60 * n,
            "gsb_scan_i32_i64": 12 * CN,
            "gsb_isect_emit": 16 * CN + 12 * I,
            "gsb_sort_pairs": (24 * 6 + 0) * I,
            "gsb_isect_offsets": 8 * I + 4 * 1200,
            "gsb_tile_scan": 12 * 1200,
            "gsb_isect_scatter": 16 * CN + 8 * I,
            "gsb_segment_sort": (8 + 8 + 4) * I,
            "gsb_composite_fwd": 52 * I + (4 * 5 + 8) * px,
            "gsb_composite_bwd": 52 * I + (8 * 5 + 12) * px + 48 * I,
            "gsb_photo_loss": (20 + 12 + 20) * px,
            "gsb_scale_by_scalar": 40 * px,
        }
        # DRAM bytes per launch from the committed `ncu --set full` capture (tools/ncu_traffic.py)
        # (tools/ncu_traffic.py stamps the file with the hash of the kernel sources it was captured from;
        # a capture of other sources is refused rather than quoted)
        traffic, traffic_note = {}, None
        tpath = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tpath) and n == 200_000:
            with open(tpath) as f:
                tj = json.load(f)
            if tj.get("csrc_sha") == csrc_sha():
                traffic = tj.get("dram_bytes_per_launch", {})
                traffic_note = f"ncu --set full capture of csrc {tj.get('csrc_sha')} ({tj.get('captured', '?')})"
            else:
                traffic_note = (f"profiles/traffic.json is of csrc {tj.get('csrc_sha')}, this tree is {csrc_sha()}: "
                                "stale capture refused, traffic = null")
        stages = []
        for name, (calls, msk) in sorted(kt.items(), key=lambda kv: -kv[1][1]):
            if name not in alg:
                continue
            ach = alg[name] / (msk * 1e-3) / 1e9
            stages.append({"kernel": name, "ms": round(msk, 5), "bound": "hbm", "achieved": round(ach, 2),
                           "peak": hbm, "unit": "GB/s", "frac": round(ach / hbm, 5),
                           "algorithmic_bytes": alg[name], "traffic": traffic.get(name)})
        # What bounds each stage.  The compositing kernels walk P (pixel, Gaussian) pairs out of
        # L2-resident lists (DRAM traffic is a tenth of their algorithmic bytes): their roofline is the
        # FP32 lanes, 148 SM x 128 lanes x 2 flop x f_SM, against the oracle-defined pair count P times
        # the flops per pair of SURVEY 8d; the HBM fraction stays beside it as `hbm_frac`.  The one-CTA
        # tile scan is a latency chain (its bytes are nothing).
        fp32_peak_tflops = 148 * 128 * 2 * sm_max * 1e6 / 1e12
        flop_per_pair = {"gsb_composite_fwd": 2 * (9 + 5) + 1, "gsb_composite_bwd": 45 + 3 * 5 + 1}
        for st in stages:
            if st["kernel"] in flop_per_pair:
                fpp = flop_per_pair[st["kernel"]]
                tfl = P * fpp / (st["ms"] * 1e-3) / 1e12
                st.update({"bound": "fp32", "hbm_frac": st["frac"], "hbm_achieved_gbs": st["achieved"],
                           "achieved": round(tfl, 3), "peak": round(fp32_peak_tflops, 2), "unit": "TFLOP/s",
                           "frac": round(tfl / fp32_peak_tflops, 5), "pairs": P, "flop_per_pair": fpp,
                           "achieved_gpairs_s": round(P / (st["ms"] * 1e-3) / 1e9, 3),
                           "peak_how": f"148 SM x 128 FP32 lanes x 2 flop x {sm_max:.0f} MHz (nominal; no measured "
                                       "FP32 figure in MEASURED_PEAKS.json)"})
            elif st["kernel"] == "gsb_tile_scan":
                st.update({"bound": "latency", "note": "one CTA, serial scan of C*T counters"})
        dom = stages[0] if stages else None
        roofline = None
        if dom:
            roofline = {"kernel": dom["kernel"], "bound": dom["bound"], "achieved": dom["achieved"],
                        "peak": dom["peak"], "unit": dom["unit"], "frac": dom["frac"], "traffic": dom["traffic"],
                        "peak_source": which if dom["bound"] == "hbm" else "nominal FP32 lanes",
                        "ms_per_launch": dom["ms"]}
            for k in ("hbm_frac", "pairs", "flop_per_pair", "algorithmic_bytes"):
                if k in dom:
                    roofline[k] = dom[k]
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload_string(n, world),
                       "l2": "256 MiB memset between timed steps, outside each step's event pair",
                       "n_isects": I, "pairs": P},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "how": "software pipeline: ONE copy of the pinned host arena -> staging on a copy stream while the "
                           "previous step computes (it starts after that step's e0 and is awaited before its e1), "
                           "staging -> step tensors D2D, graph replay, loss + pose gradient D2H to one of two pinned "
                           "result buffers; the host reads step i's result after queueing step i+1",
                    "serial_value": e2e_serial, "h2d_copy_ms": h2d_ms},
            "gpu_launches": launches,
            "launch_calls_per_step": launches / args.steps,
            "timed_loop": "CUDA-graph replay of step() (public rasterization() API + loss + backward)",
            "eager_ms_per_step": eager_ms,
            "wall_s_timed_region": wall,
            "clocks": clocks,
            "roofline": roofline,
            "traffic_source": traffic_note,
            "roofline_stages": stages,
            "mapping": mapping,
            "mapping_big": mapping_big,
            "tracking_closure": closure,
            "tracking_frames": track_frames,
            "sequence": slam_seq,
            "sequence_sharded": seq_sharded,
        }
        parity_ok = True
        if world == 1 and not args.no_cpu_baseline:
            # the oracle's step on the identical inputs: timed (cpu_baseline) AND compared (parity)
            v, tsec, cores, sample, ref = cpu_oracle_rate(n, reps=1, keep=True)
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                                    "sample": sample, "seconds": tsec}
            ref64 = None
            if not args.no_f64:
                _, ref64 = cpu_oracle_step(make_scene(n, 1, "tum", 1.0), target_images(1, H, W), torch.float64)
            line["parity"] = parity_report(gpu_step, ref, ref64)
            line["parity"]["what"] = (f"GPU step vs oracle/torch_ref.py on the identical {n}-Gaussian 640x480 view "
                                      "(relative L2 errors; images <= 1e-4, gradients <= 1e-3)")
            parity_ok = line["parity"]["ok"]
            if not args.no_cfg1:
                try:
                    line["cfg1"] = cfg1_leg(dev)
                    parity_ok = parity_ok and line["cfg1"]["parity"]["ok"]
                except Exception as e:   # noqa: BLE001  (a secondary leg must not take the headline line down)
                    line["cfg1"] = {"error": f"{type(e).__name__}: {e}",
                                    "where": traceback.format_exc().splitlines()[-12:]}
                    parity_ok = False
        partial["line"] = line
    # the 3 M-Gaussian leg last: the one that moves the most memory through the fabric cannot take the others with it
    if args.map_big == "on" or (args.map_big == "auto" and world == 8 and not args.no_mapping):
        try:
            mapping_big = mapping_big_leg(args, dev, rank, world)
        except Exception as e:   # noqa: BLE001  (reported in place; every rank fails or none: the step is collective)
            import traceback as _tb
            mapping_big = {"error": f"{type(e).__name__}: {e}", "where": _tb.format_exc().splitlines()[-8:]}
    progress('mapping_big leg done' if mapping_big is not None else 'mapping_big leg skipped')
    if rank == 0:
        # the mapping object goes LAST so that a truncated tail of the line still carries it
        line.pop("mapping_big", None)
        line["mapping_big"] = mapping_big
        line["mapping"] = line.pop("mapping")
        watchdog.cancel()
        print(json.dumps(line), flush=True)
        if not parity_ok:
            print("bench.py: PARITY FAILED (see the `parity` object of the line above)", file=sys.stderr, flush=True)
            if world == 1:
                sys.exit(3)
    watchdog.cancel()
    if world > 1:
        # Tearing the communicator down while captured graphs that hold NCCL kernels are alive can
        # block in ncclCommDestroy: drop the graphs, agree that everybody is done, and leave.
        del gstep
        import gc
        gc.collect()
        barrier()
        sys.stdout.flush()
        sys.stderr.flush()
        os._exit(0)


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
