"""Per-unit timing trace of the compositing kernels (library built with `make TRACE=1`).

    python tools/trace_units.py fwd|bwd [n_gauss]

Prints how long every unit (one warp's quarter- or half-tile) ran, when it started, how the units
were spread over the SMs and how many were alive over the kernel's span -- the numbers behind the
residency choices in composite.cu / composite_bwd.cu.
"""
import ctypes
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
from gslam_b200 import _lib  # noqa: E402
from gslam_b200.rasterization import rasterization  # noqa: E402
from gslam_b200.synthetic import make_scene  # noqa: E402

which = sys.argv[1] if len(sys.argv) > 1 else "fwd"
n_gauss = int(sys.argv[2]) if len(sys.argv) > 2 else 200_000
g = make_scene(n_gauss, 1, "tum", 1.0).to("cuda")


def run():
    ps = [q.clone().requires_grad_(True) for q in g.params()]
    out = rasterization(ps[0], ps[1], ps[2], ps[3], ps[4], g.viewmats, g.Ks, g.width, g.height, packed=False,
                        render_mode="RGB+D", log_uncertainties=ps[5], backgrounds=torch.zeros(1, 3, device="cuda"))
    out.render.sum().backward()


for _ in range(5):
    run()
torch.cuda.synchronize()
L = ctypes.CDLL(_lib.LIB_PATH)
tiles = ((g.width + 15) // 16) * ((g.height + 15) // 16)
buf = np.zeros(4 * 65536, dtype=np.uint64)      # GSB_TRACE_UNITS items of 4 words
rc = getattr(L, "gsb_debug_trace_" + which)(buf.ctypes.data_as(ctypes.c_void_p), buf.size // 4)
t = buf.reshape(-1, 4)
t = t[t[:, 1] > 0]
t = t[t[:, 0] + np.uint64(1_000_000) > t[:, 0].max()]      # the last run only (split ids of earlier runs linger)
t0 = t[:, 0].min()
beg, end = (t[:, 0] - t0).astype(np.int64), (t[:, 1] - t0).astype(np.int64)
dur, sm = end - beg, t[:, 2].astype(np.int64)
n, kept = (t[:, 3] >> np.uint64(32)).astype(np.int64), (t[:, 3] & np.uint64(0xFFFFFFFF)).astype(np.int64)
print(f"{which}: rc {rc}  units {len(t)}  kernel span {end.max()} ns")
print("unit duration ns: mean %.0f p50 %.0f p90 %.0f p99 %.0f max %.0f" % (dur.mean(), *np.percentile(dur, [50, 90, 99]), dur.max()))
print("unit start ns:    p50 %.0f p90 %.0f p99 %.0f max %.0f" % (*np.percentile(beg, [50, 90, 99]), beg.max()))
print("list length: mean %.0f max %d;  kept entries: mean %.0f max %d" % (n.mean(), n.max(), kept.mean(), kept.max()))
sms = np.unique(sm)
ends = np.array([end[sm == s].max() for s in sms])
work = np.array([kept[sm == s].sum() for s in sms])
cnt = np.array([(sm == s).sum() for s in sms])
print("per-SM finish ns: min %.0f mean %.0f max %.0f   (%d SMs)" % (ends.min(), ends.mean(), ends.max(), len(sms)))
print("per-SM kept entries: min %d mean %.0f max %d;  units per SM: min %d max %d" % (work.min(), work.mean(), work.max(), cnt.min(), cnt.max()))
print("corr(SM work, SM finish) = %.3f;  sum(unit time) / kept entries = %.1f ns" % (np.corrcoef(work, ends)[0, 1], dur.sum() / max(kept.sum(), 1)))
ts = np.linspace(0, end.max(), 21)
print("alive units at 0,5,..100 %% of the span:", [int(((beg <= x) & (end > x)).sum()) for x in ts])
mx = np.array([kept[sm == s].max() for s in sms])
A = np.stack([work, mx, np.ones_like(work)], 1).astype(np.float64)
coef, res, *_ = np.linalg.lstsq(A, ends.astype(np.float64), rcond=None)
pred = A @ coef
print("SM finish ~ %.2f ns x SM kept + %.1f ns x heaviest unit's kept + %.0f ns;  R^2 = %.3f;  corr(finish, heaviest) = %.3f"
      % (coef[0], coef[1], coef[2], 1 - ((ends - pred) ** 2).sum() / ((ends - ends.mean()) ** 2).sum(), np.corrcoef(mx, ends)[0, 1]))
# how a unit's duration depends on its own work and on its SM's load
smw = {s: w for s, w in zip(sms, work)}
B = np.stack([kept, np.array([smw[s] for s in sm]), np.ones_like(kept)], 1).astype(np.float64)
c2, *_ = np.linalg.lstsq(B, dur.astype(np.float64), rcond=None)
p2 = B @ c2
print("unit duration ~ %.1f ns x own kept + %.2f ns x SM kept + %.0f ns;  R^2 = %.3f"
      % (c2[0], c2[1], c2[2], 1 - ((dur - p2) ** 2).sum() / ((dur - dur.mean()) ** 2).sum()))
