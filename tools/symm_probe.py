"""Probe: does torch symmetric memory (CUDA backend) rendezvous on this box, and is there a multicast (NVLS) address?"""
import os, sys, torch, torch.distributed as dist
import torch.distributed._symmetric_memory as symm_mem
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
try:
    t = symm_mem.empty(1 << 20, dtype=torch.float32, device=dev)
    h = symm_mem.rendezvous(t, group=dist.group.WORLD)
    t.fill_(rank + 1)
    h.barrier()
    peer = h.get_buffer((rank + 1) % world, (8,), torch.float32)
    print(rank, "rendezvous ok; peer value", float(peer[0]), "multicast_ptr", hex(h.multicast_ptr or 0),
          "n buffer ptrs", len(h.buffer_ptrs), "signal pads", len(h.signal_pad_ptrs), flush=True)
except Exception as e:
    print(rank, "FAILED", type(e).__name__, str(e)[:400], flush=True)
dist.barrier()
dist.destroy_process_group()
