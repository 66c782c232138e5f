import cProfile, pstats, sys, time, torch
sys.path.insert(0, '.')
from gslam_b200.rasterization import rasterization
from gslam_b200.losses import tracking_loss
from gslam_b200.synthetic import make_scene, target_images
dev = torch.device("cuda:0")
g = make_scene(200_000, 1, "tum", 1.0).to(dev)
gt = target_images(1, g.height, g.width).to(dev)
ps = [p.clone().requires_grad_(True) for p in g.params()]
vm = g.viewmats.clone().requires_grad_(True)
bg = torch.zeros(1, 3, device=dev)
def step():
    for p in ps + [vm]: p.grad = None
    out = rasterization(ps[0], ps[1], ps[2], ps[3], ps[4], vm, g.Ks, g.width, g.height, packed=False,
                        render_mode="RGB+D", log_uncertainties=ps[5], backgrounds=bg)
    loss = tracking_loss(out.render, gt, None, "active-nerf")
    loss.backward()
    return loss
for _ in range(20): step()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(200): step()
torch.cuda.synchronize()
print("eager ms/step", (time.perf_counter() - t0) * 5)
pr = cProfile.Profile(); pr.enable()
for _ in range(200): step()
torch.cuda.synchronize()
pr.disable()
st = pstats.Stats(pr); st.sort_stats("tottime").print_stats(22)
