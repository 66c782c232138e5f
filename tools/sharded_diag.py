"""Diagnostic of the view-sharded sequence loop: torchrun --nproc-per-node N tools/sharded_diag.py [frames]
Runs the RGB-D sequence with the mapping window sharded over all ranks, with and without graph replay, and
prints per-run statistics plus whether every rank ends with the same map."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, ".")
from gslam_b200.slam import MapConfig, run_sequence  # noqa: E402

rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
F = int(sys.argv[1]) if len(sys.argv) > 1 else 60
modes = (True,) if len(sys.argv) > 2 and sys.argv[2] == "graph-only" else (True, False)
for graph in modes:
    slam, fps, gt = run_sequence(F, n_gauss=20_000, scale=1.0, device=dev,
                                 map_conf=MapConfig(use_gt_depths=True, deterministic=True, graph_mapping=graph),
                                 group=None)
    st = slam.stats
    est = torch.stack(slam.frame_poses)
    err = (est[:, :3, 3] - gt[:, :3, 3]).norm(dim=1)
    sig = torch.stack([slam.splats.flat.double().sum(), slam.splats.flat.double().abs().max(),
                       torch.tensor(float(slam.splats.n), device=dev, dtype=torch.float64)])
    sigs = [torch.zeros_like(sig) for _ in range(dist.get_world_size())]
    dist.all_gather(sigs, sig)
    if rank == 0:
        same = all(torch.equal(s, sigs[0]) for s in sigs)
        print(f"graph={graph} world={dist.get_world_size()} frames={F} fps={fps:.1f} keyframes={st.keyframes} n_splats={st.n_splats} "
              f"iters={st.mapping_iters} rollbacks={st.mapping_overflow_rollbacks} inserted={st.inserted} pruned={st.pruned} "
              f"err_mean={float(err.mean()):.4f} err@10,20,30,40,59={[round(float(err[min(i, F - 1)]), 4) for i in (10, 20, 30, 40, 59)]} "
              f"same_map_on_all_ranks={same} |param|max={float(sigs[0][1]):.3g}", flush=True)
    del slam
    torch.cuda.empty_cache()
dist.barrier()
torch.cuda.synchronize()
os._exit(0)
