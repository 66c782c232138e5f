"""Per-frame trace of the sequence loop (gslam_b200.slam): tracking loss / evaluations, keyframe
decisions, map size.  Usage: python tools/slam_trace.py [frames] [scale]"""
import sys

import torch

sys.path.insert(0, ".")
from gslam_b200.slam import MapConfig, Slam, synthetic_sequence      # noqa: E402
from gslam_b200.tracking import TrackingConfig                       # noqa: E402

F = int(sys.argv[1]) if len(sys.argv) > 1 else 30
scale = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
dev = torch.device("cuda:0")
cam, imgs, gt, depths = synthetic_sequence(F, 20_000, scale, dev)
rgbd = len(sys.argv) > 3 and sys.argv[3] == 'rgbd'
slam = Slam(cam, MapConfig(use_gt_depths=rgbd), TrackingConfig(), dev)
for i in range(F):
    kf0, it0, ev0 = slam.stats.keyframes, slam.stats.mapping_iters, slam.stats.tracking_evals
    tt0, tm0 = slam.stats.t_tracking, slam.stats.t_mapping
    vm = slam.process_frame(imgs[i], 1, depths[i])
    st = slam.stats
    rel = (vm @ torch.linalg.inv(gt[i]))
    print(f"frame {i:3d} track_loss {st.last_tracking_loss:.5f} evals {st.tracking_evals - ev0:3d} "
          f"kf {'+' if st.keyframes > kf0 else ' '}{st.keyframes:3d} map_iters +{st.mapping_iters - it0:3d} "
          f"photo {st.last_photometric:.4f} N {st.n_splats:5d} paused {int(slam.pause_map_optim)} "
          f"|t_est| {float(vm[:3, 3].norm()):.4f} |t_gt| {float(gt[i][:3, 3].norm()):.4f} "
          f"rot_err {float((rel[:3, :3] - torch.eye(3, device=dev)).norm()):.4f} "
          f"track {1e3 * (st.t_tracking - tt0):6.1f} ms map {1e3 * (st.t_mapping - tm0):6.1f} ms captures {slam.tracker.n_captures}")
print(st)
