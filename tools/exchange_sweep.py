"""Per-level timing of the distributed build's fused paths (BASELINE.json config 5: "per-level exchange GB/s
sweep against 900 GB/s per direction").  torchrun --nproc-per-node W tools/exchange_sweep.py [log2n] [dim]

For every distributed level: device time of the select (sample bracket, digit passes over the candidates, the
histograms reduced through NVLS multimem.red or one-shot P2P slots, scans, count exchange) and
of the partition-to-peer kernel, the bytes that kernel pushes over NVLink from this GPU, and the resulting
GB/s per direction per GPU.  Then the local phase (whose stores are replicated into every GPU = the all-gather).
Times are max over ranks (CUDA events)."""
import ctypes as C
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

import parallel_kd_tree_b200 as kd
from parallel_kd_tree_b200.parallel import dist_build as db

rank, world, local_rank = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local_rank)
dev = torch.device("cuda", local_rank)
os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
dist.init_process_group("nccl", device_id=dev)
log2n = int(sys.argv[1]) if len(sys.argv) > 1 else 24
dim = int(sys.argv[2]) if len(sys.argv) > 2 else 8
n = 1 << log2n
first, local = kd.models.slice_bounds(n, world, rank)
dist_name = sys.argv[3] if len(sys.argv) > 3 else "uniform"
if dist_name == "clustered":   # Gaussian blobs (BASELINE.json config 4): skewed medians, uneven redistribution
    pts = kd.ops.generate_clustered(1, dim, n)[first:first + local].contiguous().to(dev)
else:
    pts = kd.ops.generate_uniform(1, dim, local, first).to(dev)
b = db.IpcDistBuild(n, dim, with_tree_points=True)
L = kd.lib()
st = torch.cuda.current_stream().cuda_stream
levels = world.bit_length() - 1
steps = db.level_schedule()


def ev():
    return torch.cuda.Event(enable_timing=True)


def run(record):
    b.serial += 1
    b.load(pts)
    for level in range(levels):
        e0, e1, e2, e3 = ev(), ev(), ev(), ev()
        e0.record()
        for ph, ps in steps:       # select = everything before the partition kernel; sync/fix-up = everything after it
            if ph == db.PH_PART:
                e1.record()
            kd._lib.check(L.kd_dist_phase(C.byref(b.ctx), level, ph, ps, b.serial, st), "phase")
            if ph == db.PH_PART:
                e2.record()
        e3.record()
        record.append((level, e0, e1, e2, e3))
    e4, e5 = ev(), ev()
    e4.record()
    kd._lib.check(L.kd_dist_local_build(C.byref(b.ctx), st), "local")
    kd._lib.check(L.kd_dist_done(C.byref(b.ctx), 1, b.serial, st), "done")
    kd._lib.check(L.kd_dist_done(C.byref(b.ctx), 0, b.serial, st), "done")
    e5.record()
    record.append(("local", e4, e5))


run([])
torch.cuda.synchronize(); dist.barrier()
rec = []
run(rec)
torch.cuda.synchronize()
b.status()
rows = []
for r in rec:
    if r[0] == "local":
        t = torch.tensor([r[1].elapsed_time(r[2])], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        nloc = n // world
        out_bytes = nloc * 4 * (1 + dim) * (world - 1)       # perm + tree_points slices stored into every other GPU
        rows.append({"phase": "local build + fused all-gather", "ms": round(float(t), 3),
                     "nvlink_bytes_out_per_gpu": out_bytes,
                     "note": "lower bound of the replication alone at 770 GB/s: %.3f ms" % (out_bytes / 770e9 * 1e3)})
        continue
    level, e0, e1, e2, e3 = r
    G = world >> level
    t = torch.tensor([e0.elapsed_time(e1), e1.elapsed_time(e2), e2.elapsed_time(e3)], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    nloc = n // world
    out_bytes = int(nloc * (1 - 1 / G) * 4 * (dim + 1))      # expected share of my points whose owner is another GPU
    gbs = out_bytes / (float(t[1]) * 1e-3) / 1e9
    rows.append({"level": level, "group": G, "select_ms": round(float(t[0]), 3), "partition_ms": round(float(t[1]), 3),
                 "sync_fixup_ms": round(float(t[2]), 3), "nvlink_bytes_out_per_gpu": out_bytes,
                 "exchange_GBps_per_direction": round(gbs, 1), "frac_of_770_measured": round(gbs / 770, 3),
                 "frac_of_900_nominal": round(gbs / 900, 3)})
# Load-balance report: bytes every rank pushes to every peer per level, from the ranks' own
# counters (an extra, untimed build that stops after every level to read them).
def owner_spans(first, count, total, members):
    """[(member, n)] of the output slots [first, first+count) of a child range of `total` slots spread over `members`."""
    base, rem, out, o = total // members, total % members, [], first
    while count > 0:
        big = rem * (base + 1)
        mem = o // (base + 1) if o < big else rem + (o - big) // base
        start = mem * (base + 1) if mem < rem else big + (mem - rem) * base
        size = base + 1 if mem < rem else base
        k = min(count, start + size - o)
        out.append((mem, k))
        o += k
        count -= k
    return out


def seg(N, level, s):
    lo, cnt = 0, N
    for b in range(level - 1, -1, -1):
        m = cnt // 2
        if (s >> b) & 1:
            lo, cnt = lo + m + 1, cnt - m - 1
        else:
            cnt = m
    return lo, cnt


balance = []
b.serial += 1
b.load(pts)
for level in range(levels):
    for ph, ps in steps:
        kd._lib.check(L.kd_dist_phase(C.byref(b.ctx), level, ph, ps, b.serial, st), "phase")
    cnt = (C.c_uint32 * 8)()
    kd._lib.check(L.kd_dist_level_counts(C.byref(b.ctx), cnt, st), "counts")
    G = world >> level
    gb, j = (rank // G) * G, rank % G
    _, gn = seg(n, level, rank // G)
    m, nr, Gh = gn // 2, gn - gn // 2 - 1, G // 2
    local_n, lc, mine, Loff, Roff = cnt[0], cnt[1], cnt[2], cnt[3], cnt[4]
    free_n = local_n - 1 if j == G - 1 else local_n
    rc = free_n - lc - mine
    row = [0] * world
    for mem, k in owner_spans(Loff, lc, m, Gh):
        row[gb + mem] += k * 4 * (dim + 1)
    for mem, k in owner_spans(Roff, rc, nr, Gh):
        row[gb + Gh + mem] += k * 4 * (dim + 1)
    allrows = [None] * world
    dist.all_gather_object(allrows, {"rank": rank, "bytes_to_peer": row, "bracket_used": int(cnt[7]),
                                     "candidates": int(cnt[6]), "local_points": int(local_n)})
    if rank == 0:
        sent = [sum(v for p, v in enumerate(r["bytes_to_peer"]) if p != r["rank"]) for r in allrows]
        recv = [sum(r["bytes_to_peer"][p] for r in allrows if r["rank"] != p) for p in range(world)]
        balance.append({"level": level, "ranks": allrows, "max_over_mean_sent": round(max(sent) / (sum(sent) / world), 3),
                        "max_over_mean_received": round(max(recv) / (sum(recv) / world), 3)})
kd._lib.check(L.kd_dist_local_build(C.byref(b.ctx), st), "local")
kd._lib.check(L.kd_dist_done(C.byref(b.ctx), 1, b.serial, st), "done")
kd._lib.check(L.kd_dist_done(C.byref(b.ctx), 0, b.serial, st), "done")
kd._lib.check(L.kd_dist_finish(C.byref(b.ctx), st), "finish")
torch.cuda.synchronize()
b.status()
if rank == 0:
    # link bytes of a level = what the busiest GPU really pushed to other GPUs (from the counters above), not an estimate
    for r in rows:
        if "level" not in r:
            continue
        bl = balance[r["level"]]
        sent = [sum(v for p, v in enumerate(x["bytes_to_peer"]) if p != x["rank"]) for x in bl["ranks"]]
        recv = [sum(x["bytes_to_peer"][p] for x in bl["ranks"] if x["rank"] != p) for p in range(world)]
        busiest = max(max(sent), max(recv))
        gbs = busiest / (r["partition_ms"] * 1e-3) / 1e9
        r.update({"nvlink_bytes_busiest_gpu": busiest, "nvlink_bytes_mean_per_gpu": int(sum(sent) / world),
                  "exchange_GBps_per_direction": round(gbs, 1), "frac_of_770_measured": round(gbs / 770, 3),
                  "frac_of_900_nominal": round(gbs / 900, 3)})
        r.pop("nvlink_bytes_out_per_gpu", None)
    print(json.dumps({"n": n, "dim": dim, "world": world, "distribution": dist_name, "rows": rows,
                      "per_peer_bytes_per_level": balance}, indent=1))
b.close()
dist.destroy_process_group()
