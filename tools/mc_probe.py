"""Does torch symmetric memory hand out an NVLS multicast mapping on this box?  (torchrun, >= 2 ranks)"""
import os
import torch
import torch.distributed as dist
import torch.distributed._symmetric_memory as symm

rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
t = symm.empty(1 << 20, dtype=torch.float32, device=dev)
h = symm.rendezvous(t, dist.group.WORLD)
print(rank, "multicast_ptr", hex(int(getattr(h, "multicast_ptr", 0) or 0)), "has_multicast_support",
      getattr(symm, "has_multicast_support", lambda *a: None)(dev.type, local) if hasattr(symm, "has_multicast_support") else None, flush=True)
dist.destroy_process_group()
