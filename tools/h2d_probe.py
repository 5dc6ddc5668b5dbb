"""Host<->device copy bandwidth per rank, alone and with all ranks copying at once (end-to-end scaling diagnosis).
    python -m torch.distributed.run --nproc-per-node W tools/h2d_probe.py [MiB]
Pinned host buffers; CUDA events; every rank reports GB/s for H2D and D2H (a) while the others idle, (b) all together,
(c) all together with the process pinned to the CPUs of the GPU's NUMA node (from `nvidia-smi topo -C`-style sysfs)."""
import json
import os
import sys
import time

import torch
import torch.distributed as dist

rank, world, local_rank = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local_rank)
dev = torch.device("cuda", local_rank)
os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
dist.init_process_group("nccl", device_id=dev)
mib = int(sys.argv[1]) if len(sys.argv) > 1 else 384
nbytes = mib << 20


def numa_node_of_gpu():
    try:
        bus = torch.cuda.get_device_properties(dev).pci_bus_id
        dom = torch.cuda.get_device_properties(dev).pci_domain_id
        devid = torch.cuda.get_device_properties(dev).pci_device_id
        path = f"/sys/bus/pci/devices/{dom:04x}:{bus:02x}:{devid:02x}.0/numa_node"
        return int(open(path).read())
    except Exception:
        return -1


def bw(h, d, to_dev, reps=5):
    best = 0.0
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        (d.copy_(h, non_blocking=True) if to_dev else h.copy_(d, non_blocking=True))
        b.record()
        torch.cuda.synchronize()
        best = max(best, nbytes / (a.elapsed_time(b) * 1e-3) / 1e9)
    return best


host = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
devb = torch.empty(nbytes, dtype=torch.uint8, device=dev)
bw(host, devb, True, 2)
res = {"rank": rank, "numa_node": numa_node_of_gpu(), "cpus_allowed": len(os.sched_getaffinity(0))}
alone = [0.0, 0.0]
for r in range(world):
    dist.barrier()
    if r == rank:
        alone = [bw(host, devb, True), bw(host, devb, False)]
dist.barrier()
together = [bw(host, devb, True), 0.0]
dist.barrier()
together[1] = bw(host, devb, False)
dist.barrier()
# both directions at once on two streams (what the pipelined e2e loop does)
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
host2 = torch.empty(nbytes // 3, dtype=torch.uint8, pin_memory=True)
dev2 = torch.empty(nbytes // 3, dtype=torch.uint8, device=dev)
torch.cuda.synchronize(); dist.barrier()
t0 = time.perf_counter()
for _ in range(5):
    with torch.cuda.stream(s1):
        devb.copy_(host, non_blocking=True)
    with torch.cuda.stream(s2):
        host2.copy_(dev2, non_blocking=True)
torch.cuda.synchronize()
duplex_ms = (time.perf_counter() - t0) / 5 * 1e3
res.update({"h2d_alone_GBps": round(alone[0], 1), "d2h_alone_GBps": round(alone[1], 1),
            "h2d_all_ranks_GBps": round(together[0], 1), "d2h_all_ranks_GBps": round(together[1], 1),
            "duplex_h2d_%dMiB_plus_d2h_%dMiB_all_ranks_ms" % (mib, mib // 3): round(duplex_ms, 2)})
out = [None] * world
dist.all_gather_object(out, res)
if rank == 0:
    print(json.dumps({"mib": mib, "world": world, "ranks": out}))
dist.destroy_process_group()
