"""Event-timed gsb_project_fwd / gsb_project_bwd pieces: with and without tile counting, by size."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gslam_b200 import ops
from gslam_b200.synthetic import make_scene
dev = torch.device("cuda:0")
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

def timeit(fn, reps=20):
    for _ in range(3):
        fn()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
    for a, b in ev:
        flush.zero_()
        torch.cuda._sleep(200_000)
        a.record(); fn(); b.record()
    torch.cuda.synchronize()
    ts = sorted(x.elapsed_time(y) for x, y in ev)
    return ts[len(ts) // 2] * 1e3

for n, c in ((200_000, 1), (500_000, 1), (500_000, 8)):
    sc = make_scene(n, c, "tum", 1.0).to(dev)
    P = sc.params()
    th, tw = 30, 40
    counts = torch.zeros(c, th * tw, 32, dtype=torch.int32, device=dev)
    def fwd(tc):
        with torch.no_grad():
            return ops._Project.apply(P[0], P[1], P[2], P[3], P[4], P[5], sc.viewmats, sc.Ks, sc.width, sc.height,
                                      0.3, 0.01, 1e10, 0.0, tw, th, tc, None, False)
    t_with = timeit(lambda: fwd(counts))
    # without counting: pass a NULL tile_counts through the raw call
    from gslam_b200._lib import call, ptr, stream
    radii = torch.empty(c, n, dtype=torch.int32, device=dev); tiles = torch.empty_like(radii)
    m2 = torch.empty(c, n, 2, device=dev); rec = torch.empty(c, n, 12, device=dev)
    def raw(tc):
        call("gsb_project_fwd", ptr(P[0]), ptr(P[1]), ptr(P[2]), ptr(P[3]), ptr(P[4]), ptr(P[5]), ptr(sc.viewmats), ptr(sc.Ks),
             c, n, sc.width, sc.height, 0.3, 0.01, 1e10, 0.0, tw, th, ptr(radii), ptr(m2), ptr(rec), ptr(tiles), tc, None, -1, 0, stream())
    t_raw_with = timeit(lambda: raw(ptr(counts)))
    t_raw_without = timeit(lambda: raw(None))
    I = int(tiles.sum())
    print(f"N={n} C={c} I={I}: project_fwd op {t_with:.1f} us, raw with counts {t_raw_with:.1f}, raw without {t_raw_without:.1f}; "
          f"HBM floor {(60*n + 64*c*n) / 6.5565e6:.1f} us")
