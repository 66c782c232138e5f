"""Per-entry-point CUDA-event times of one MappingStep (cfg3 shape) on this rank's views."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gslam_b200 import _lib
from gslam_b200.mapping import DEFAULT_LRS, PARAM_NAMES, FlatSplats, MappingStep
from gslam_b200.optim import FusedAdam
from gslam_b200.synthetic import make_scene, target_images

n_views = int(sys.argv[1]) if len(sys.argv) > 1 else 8
n = int(sys.argv[2]) if len(sys.argv) > 2 else 500_000
dev = torch.device("cuda:0")
sc = make_scene(n, n_views, "tum", 1.0)
g = sc.to(dev)
splats = FlatSplats(g.params())
opt = FusedAdam([dict(params=[p], lr=DEFAULT_LRS[nm]) for p, nm in zip(splats.params, PARAM_NAMES)], capturable=True)
step = MappingStep(splats, opt)
gt = target_images(n_views, sc.height, sc.width).to(dev)
exposure = torch.zeros(n_views, 2, device=dev)
for _ in range(3):
    step(g.viewmats, g.Ks, gt, exposure, sc.width, sc.height, n_views)
torch.cuda.synchronize()
t0 = time.perf_counter()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    step(g.viewmats, g.Ks, gt, exposure, sc.width, sc.height, n_views)
e1.record()
torch.cuda.synchronize()
print(f"step: {e0.elapsed_time(e1) / 10:.3f} ms device, {(time.perf_counter() - t0) * 100:.3f} ms wall")
_lib.start_timing()
for _ in range(5):
    step(g.viewmats, g.Ks, gt, exposure, sc.width, sc.height, n_views)
kt = _lib.stop_timing()
tot = 0.0
for name, (calls, ms) in sorted(kt.items(), key=lambda kv: -kv[1][1] * kv[1][0]):
    print(f"{name:24s} calls/step {calls / 5:4.1f}  {ms * 1e3:9.1f} us/call")
    tot += ms * calls / 5
print(f"sum of C-ABI calls: {tot:.3f} ms/step")
# the same step as ONE CUDA graph (what bench.py times): device ms per replay
from gslam_b200.graph import GraphedStep
gs = GraphedStep(lambda: step(g.viewmats, g.Ks, gt, exposure, sc.width, sc.height, n_views).total)
for _ in range(3):
    gs.replay()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20):
    gs.replay()
e1.record()
torch.cuda.synchronize()
print(f"graph replay: {e0.elapsed_time(e1) / 20:.3f} ms per step ({gs.launches_per_replay} C-ABI launches)")
