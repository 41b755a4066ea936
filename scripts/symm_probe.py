"""Probe: torch symmetric memory between the ranks of one node (peer pointers for the fused compute + all-gather path)."""
import os
import torch
import torch.distributed as dist
import torch.distributed._symmetric_memory as symm

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
t = symm.empty((world, 1 << 20), dtype=torch.float32, device=dev)
t.zero_()
h = symm.rendezvous(t, dist.group.WORLD)
print(rank, "ptrs", [hex(p) for p in h.buffer_ptrs], "rank", h.rank, "world", h.world_size, "multicast_ptr",
      hex(getattr(h, "multicast_ptr", 0) or 0), flush=True)
h.barrier(channel=0)
for p in range(world):
    peer = h.get_buffer(p, (world, 1 << 20), torch.float32)
    peer[rank].fill_(float(rank + 1))
torch.cuda.synchronize()
h.barrier(channel=0)
torch.cuda.synchronize()
print(rank, "row values", [float(t[r, 12345]) for r in range(world)], flush=True)
dist.destroy_process_group()
