import os, torch, torch.distributed as dist
rank=int(os.environ["RANK"]); torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
import torch.distributed._symmetric_memory as sm
print(rank, [x for x in dir(sm) if not x.startswith("_")][:40], flush=True)
t = sm.empty(1<<20, dtype=torch.uint8, device="cuda")
h = sm.rendezvous(t, dist.group.WORLD.group_name)
print(rank, "buffer_ptrs", [hex(p) for p in h.buffer_ptrs], "multicast", hex(h.multicast_ptr) if h.multicast_ptr else None, "signal", [hex(p) for p in h.signal_pad_ptrs][:2], flush=True)
dist.barrier(); dist.destroy_process_group()
