import sys, time
sys.path.insert(0, ".")
import torch
import parallel_kd_tree_b200 as kd
for log2n in (22, 25, 28):
    n = 1 << log2n
    pts = kd.ops.generate_uniform(1, 3, n).cuda()
    perm = torch.empty(n, dtype=torch.int32, device="cuda")
    tp = torch.empty_like(pts)
    ws = torch.empty(kd.ops.build_workspace_bytes(n, 3), dtype=torch.uint8, device="cuda")
    for _ in range(3):
        kd.ops.build_into(pts, perm, tp, ws)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    a.record()
    kd.ops.build_into(pts, perm, tp, ws)
    b.record()
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    print(f"2^{log2n}: host launch time {1e3*(t1-t0):.3f} ms, device {a.elapsed_time(b):.3f} ms, launches {kd.ops.last_launch_count()}")
    # graph replay
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        kd.ops.build_into(pts, perm, tp, ws)
        torch.cuda.synchronize()
        with torch.cuda.graph(g, stream=s):
            kd.ops.build_into(pts, perm, tp, ws)
    torch.cuda.synchronize()
    ref = perm.clone()
    perm.fill_(-1)
    for _ in range(2):
        g.replay()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    a.record(); g.replay(); b.record()
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    print(f"   graph: host {1e3*(t1-t0):.3f} ms, device {a.elapsed_time(b):.3f} ms, same result {torch.equal(perm, ref)}")
