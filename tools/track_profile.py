"""Where a tracked frame's wall time goes (gslam_b200.tracking.Tracker.track_lbfgs): host-side
phases timed with perf_counter around explicit syncs.  Usage: python tools/track_profile.py [N]"""
import sys
import time

import torch

sys.path.insert(0, ".")
from gslam_b200 import lbfgs as L                                    # noqa: E402
from gslam_b200.map import GaussianSplattingData                     # noqa: E402
from gslam_b200.primitives import Camera                             # noqa: E402
from gslam_b200.synthetic import make_scene, smooth_trajectory       # noqa: E402
from gslam_b200.tracking import Tracker, TrackingConfig, constant_motion  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 200_000
dev = torch.device("cuda:0")
sc = make_scene(n, 1, "tum", 1.0).to(dev)
splats = GaussianSplattingData(*sc.params(), torch.zeros(n, dtype=torch.long, device=dev)).no_grad_clone()
cam = Camera(sc.Ks[0].clone(), sc.height, sc.width)
poses = smooth_trajectory(12).to(dev)
with torch.no_grad():
    imgs = [splats.render([cam], [poses[i]]).rgbs[0].clamp(0, 1).clone() for i in range(12)]
tr = Tracker(cam, TrackingConfig(), dev)
tr.set_map(splats)
tr.track_lbfgs(imgs[1], poses[0])
sync = torch.cuda.synchronize
sync()

# 1. raw graph replays
for kind in ("adam", "closure"):
    g = tr._graphs[kind]
    sync(); t0 = time.perf_counter()
    for _ in range(50):
        g.replay()
    sync(); t1 = time.perf_counter()
    print(f"{kind:8s} graph: {1e3 * (t1 - t0) / 50:.3f} ms per replay (50 back to back), "
          f"{g.launches_per_replay} C-ABI launches")
g = tr._graphs["closure"]
st = torch.cuda.current_stream()
sync(); t0 = time.perf_counter()
for _ in range(50):
    g.replay(); st.synchronize()
t1 = time.perf_counter()
print(f"closure replay + stream sync: {1e3 * (t1 - t0) / 50:.3f} ms")

# 2. phases of one frame
orig_min = L.lbfgs_minimize
acc = {"eval": 0.0, "n": 0}
import gslam_b200.tracking as T
def timed_min(evaluate, *a, **k):
    def ev(x):
        t = time.perf_counter(); r = evaluate(x); acc["eval"] += time.perf_counter() - t; acc["n"] += 1
        return r
    return orig_min(ev, *a, **k)
T.lbfgs_minimize = timed_min
for i in range(2, 8):
    acc["eval"], acc["n"] = 0.0, 0
    sync(); t0 = time.perf_counter()
    init = constant_motion(poses[i - 2], poses[i - 1]); sync()
    t1 = time.perf_counter()
    tr._begin(imgs[i], init, None, None); sync()
    t2 = time.perf_counter()
    r = tr.track_lbfgs(imgs[i], init); sync()
    t3 = time.perf_counter()
    print(f"frame {i}: constant_motion {1e3 * (t1 - t0):.2f} ms | begin {1e3 * (t2 - t1):.2f} ms | track {1e3 * (t3 - t2):.2f} ms "
          f"(evals {r.n_evals}, lbfgs evals {acc['n']} taking {1e3 * acc['eval']:.2f} ms = "
          f"{1e3 * acc['eval'] / max(acc['n'], 1):.3f} ms each)")
