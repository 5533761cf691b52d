"""Probe torch symmetric memory between ranks (dev tool): allocate, rendezvous, print peer pointers, check that a
peer-written value is visible."""
import os, torch, torch.distributed as dist
import torch.distributed._symmetric_memory as symm
rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); lr = int(os.environ.get("LOCAL_RANK", rank))
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
t = symm.empty(1024, dtype=torch.float64, device=torch.device("cuda", lr))
h = symm.rendezvous(t, dist.group.WORLD)
print(rank, "rank/world", h.rank, h.world_size, "buffer_ptrs", [hex(p) for p in h.buffer_ptrs], "signal_pad_ptrs", [hex(p) for p in h.signal_pad_ptrs], "pad size", getattr(h, "signal_pad_size", None), flush=True)
t.fill_(float(rank + 1))
h.barrier()
peer = h.get_buffer((rank + 1) % world, (1024,), torch.float64)
print(rank, "peer value", float(peer[0]), flush=True)
h.barrier()
dist.destroy_process_group()
