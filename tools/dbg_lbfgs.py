import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from gslam_b200.map import GaussianSplattingData
from gslam_b200.primitives import Camera
from gslam_b200.synthetic import make_scene, smooth_trajectory
from gslam_b200.tracking import Tracker, TrackingConfig
from gslam_b200 import lbfgs as L
cuda = torch.device("cuda:0")
n = 8000
sc = make_scene(n, 1, "tum", 0.5).to(cuda)
splats = GaussianSplattingData(*sc.params(), torch.zeros(n, dtype=torch.long, device=cuda)).no_grad_clone()
cam = Camera(sc.Ks[0].clone(), sc.height, sc.width)
true = smooth_trajectory(11)[10].to(cuda)
with torch.no_grad():
    gt = splats.render([cam], [true]).rgbs[0].clamp(0, 1).clone()
# host path, recording the points
tr = Tracker(cam, TrackingConfig(deterministic=True, lbfgs_on_device=False), cuda); tr.set_map(splats)
pts_h = []
orig = L.lbfgs_minimize
def rec(evaluate, x0, **kw):
    def ev(x):
        f, g = evaluate(x); pts_h.append((np.asarray(x, dtype=np.float32).copy(), f, np.array(g).copy())); return f, g
    return orig(ev, x0, **kw)
import gslam_b200.tracking as T
T.lbfgs_minimize = rec
rh = tr.track_lbfgs(gt, torch.eye(4, device=cuda))
# device path step by step
tr2 = Tracker(cam, TrackingConfig(deterministic=True, lbfgs_on_device=True), cuda); tr2.set_map(splats)
tr2._begin(gt, torch.eye(4, device=cuda), None, None)
adam = tr2._graph("adam")
for _ in range(10): adam.replay()
g = tr2._graph("closure_dev")
tr2.dev_lbfgs.reset(tr2.theta)
pts_d = []
for k in range(27):
    th = tr2.theta.detach().cpu().numpy().copy()
    g.replay(); torch.cuda.synchronize()
    r = tr2.result.cpu().numpy().copy()
    pts_d.append((th, float(r[0]), r[1:], tr2.dev_lbfgs.summary.cpu().numpy().copy()))
for k in range(min(len(pts_h), len(pts_d))):
    dx = np.abs(pts_h[k][0] - pts_d[k][0]).max()
    print(k, "dx %.3e" % dx, "f_host %.9f f_dev %.9f" % (pts_h[k][1], pts_d[k][1]), "dg %.3e" % np.abs(pts_h[k][2] - pts_d[k][2]).max(), pts_d[k][3][:3])
print(len(pts_h), len(pts_d))
