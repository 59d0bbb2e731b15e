"""Probe: torch symmetric memory + NVLS multicast on this box (run under torchrun, 2+ ranks)."""
import datetime
import os
import sys

import torch
import torch.distributed as dist
import torch.distributed._symmetric_memory as symm

rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); local = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local), timeout=datetime.timedelta(seconds=60))
t = symm.empty(world * 1024, dtype=torch.int32, device=torch.device("cuda", local))
t.fill_(-1)
hdl = symm.rendezvous(t, dist.group.WORLD)
print(rank, "multicast_ptr", hex(hdl.multicast_ptr or 0), "buffer_ptrs", [hex(p) for p in hdl.buffer_ptrs], flush=True)
hdl.barrier(channel=0)
# every rank writes its block into every peer's buffer through the peer pointers (P2P stores from a torch kernel)
for p in range(world):
    peer = hdl.get_buffer(p, (world * 1024,), torch.int32)
    peer[rank * 1024:(rank + 1) * 1024] = rank
hdl.barrier(channel=1)
torch.cuda.synchronize()
print(rank, "blocks", t.view(world, 1024)[:, 0].tolist(), flush=True)
dist.barrier()
dist.destroy_process_group()
