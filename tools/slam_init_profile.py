"""Cost of the start of a sequence (gslam_b200.slam): initialisation, mapping iterations on the small
first map, and the per-entry-point times of those iterations."""
import sys
import time

import torch

sys.path.insert(0, ".")
from gslam_b200 import _lib                                           # noqa: E402
from gslam_b200.slam import MapConfig, Slam, synthetic_sequence       # noqa: E402

dev = torch.device("cuda:0")
cam, imgs, gt, depths = synthetic_sequence(4, 20000, 1.0, dev)
slam = Slam(cam, MapConfig(), None, dev)
torch.cuda.synchronize(); t0 = time.perf_counter()
slam.initialize(imgs[0]); torch.cuda.synchronize(); t1 = time.perf_counter()
print("initialize (insertion, knn):", round(1e3 * (t1 - t0), 1), "ms")
for k in range(3):
    it0 = slam.stats.mapping_iters
    t1 = time.perf_counter(); slam.optimize_map(50, prune=False); torch.cuda.synchronize(); t2 = time.perf_counter()
    n = slam.stats.mapping_iters - it0
    print("optimize_map:", n, "iters", round(1e3 * (t2 - t1) / max(n, 1), 3), "ms per iteration")
_lib.start_timing()
slam.optimize_map(5, prune=False)
kt = _lib.stop_timing()
print({k: (c, round(ms * 1e3, 1)) for k, (c, ms) in kt.items()})
slam.sync()
r = slam.tracker.track_lbfgs(imgs[1], torch.eye(4, device=dev))
torch.cuda.synchronize(); t0 = time.perf_counter()
r = slam.tracker.track_lbfgs(imgs[1], torch.eye(4, device=dev))
torch.cuda.synchronize(); t1 = time.perf_counter()
print(f"track: {1e3 * (t1 - t0):.2f} ms for {r.n_evals} evaluations on a {slam.splats.n}-splat map")
