"""Multi-GPU check (torchrun): the fused NVLS mapping update against the NCCL reduce-scatter / all-gather form.

Same scene, same init, deterministic accumulation, 3 steps each: parameters must be bit-identical ACROSS RANKS in
both forms, and the two forms must agree to rounding (the in-switch sum orders the ranks differently)."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
from gslam_b200 import ops
from gslam_b200.mapping import DEFAULT_LRS, PARAM_NAMES, FlatSplats, MappingStep, shard_views
from gslam_b200.optim import FusedAdam
from gslam_b200.synthetic import make_scene, target_images

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
n, n_views = int(sys.argv[1]) if len(sys.argv) > 1 else 100_003, 8
sc = make_scene(n, n_views, "tum", 0.5)
g = sc.to(dev)
mine = shard_views(n_views, rank, world)
gt = target_images(n_views, sc.height, sc.width)[mine].to(dev)
ex = torch.zeros(len(mine), 2, device=dev)
out = {}
with ops.deterministic_mode(True):
    for comm in ("nccl", "nvls"):
        sp = FlatSplats(g.params(), align=4 * world, symmetric=dist.group.WORLD if comm == "nvls" else None)
        opt = FusedAdam([dict(params=[p], lr=DEFAULT_LRS[nm]) for p, nm in zip(sp.params, PARAM_NAMES)], capturable=True)
        step = MappingStep(sp, opt, comm=comm)
        losses = []
        for _ in range(3):
            r = step(g.viewmats[mine].contiguous(), g.Ks[mine].contiguous(), gt, ex, sc.width, sc.height, n_views)
            losses.append(float(r.total))
        torch.cuda.synchronize()
        flat = torch.cat([p.detach().reshape(-1) for p in sp.params])
        gathered = [torch.empty_like(flat) for _ in range(world)]
        dist.all_gather(gathered, flat)
        same = all(torch.equal(gathered[0], x) for x in gathered)
        out[comm] = dict(flat=flat, losses=losses, same_across_ranks=same, vis=r.vis_count.clone(), nvls=step._nvls is not None)
p0 = torch.cat([p.reshape(-1) for p in g.params()])
d_nccl, d_nvls = out["nccl"]["flat"] - p0, out["nvls"]["flat"] - p0
rep = {"world": world, "n": n, "nvls_used": out["nvls"]["nvls"], "same_across_ranks": {k: v["same_across_ranks"] for k, v in out.items()},
       "displacement_rel_diff": float((d_nvls - d_nccl).norm() / d_nccl.norm()),
       "max_abs_param_diff": float((out["nvls"]["flat"] - out["nccl"]["flat"]).abs().max()),
       "losses": {k: v["losses"] for k, v in out.items()}, "vis_equal": bool(torch.equal(out["nccl"]["vis"], out["nvls"]["vis"]))}
ok = rep["nvls_used"] and all(rep["same_across_ranks"].values()) and rep["displacement_rel_diff"] < 2e-3 and rep["vis_equal"] \
    and all(abs(a - b) < 1e-5 * abs(b) for a, b in zip(rep["losses"]["nvls"], rep["losses"]["nccl"]))
rep["ok"] = bool(ok)
if rank == 0:
    print(json.dumps(rep), flush=True)
dist.barrier()
dist.destroy_process_group()
sys.exit(0 if ok else 1)
