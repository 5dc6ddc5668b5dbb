"""Where does the end-to-end step time go on ONE GPU?  Steady-state cycle of upload + build + download with the copies
(a) as in bench.py, (b) without the download, (c) without the upload, (d) upload split over two streams, (e) copies only."""
import sys
import time

sys.path.insert(0, ".")
import torch

import parallel_kd_tree_b200 as kd

n, dim = 1 << 28, 3
host_pts = kd.ops.generate_uniform(1, dim, n, pin_memory=True)
pts2 = [torch.empty((n, dim), device="cuda") for _ in range(2)]
perm2 = [torch.empty(n, dtype=torch.int32, device="cuda") for _ in range(2)]
host_perm2 = [torch.empty(n, dtype=torch.int32, pin_memory=True) for _ in range(2)]
tp = torch.empty((n, dim), device="cuda")
ws = torch.empty(kd.ops.build_workspace_bytes(n, dim), dtype=torch.uint8, device="cuda")
s_up, s_up2, s_dn = torch.cuda.Stream(), torch.cuda.Stream(), torch.cuda.Stream()
main = torch.cuda.current_stream()
pts2[0].copy_(host_pts); pts2[1].copy_(host_pts)
half = n // 2


def run(k, up=True, down=True, build=True, split=False):
    ev_build, ev_dn = [None] * k, [None] * k
    for i in range(k):
        b = i & 1
        if up:
            for s, sl in ((s_up, slice(0, half if split else n)),) + (((s_up2, slice(half, n)),) if split else ()):
                with torch.cuda.stream(s):
                    if i >= 2:
                        s.wait_event(ev_build[i - 2])
                    pts2[b][sl].copy_(host_pts[sl], non_blocking=True)
                    e = torch.cuda.Event(); e.record(s)
                main.wait_event(e)
        if i >= 2 and down:
            main.wait_event(ev_dn[i - 2])
        if build:
            kd.ops.build_into(pts2[b], perm2[b], tp, ws)
        ev_build[i] = torch.cuda.Event(); ev_build[i].record(main)
        if down:
            with torch.cuda.stream(s_dn):
                s_dn.wait_event(ev_build[i])
                host_perm2[b].copy_(perm2[b], non_blocking=True)
                ev_dn[i] = torch.cuda.Event(); ev_dn[i].record(s_dn)
    torch.cuda.synchronize()


for name, kw in [("as bench.py", {}), ("no download", {"down": False}), ("no upload", {"up": False}),
                 ("upload on two streams", {"split": True}), ("copies only", {"build": False}),
                 ("upload only", {"build": False, "down": False}), ("build only", {"up": False, "down": False})]:
    run(2, **kw)
    t0 = time.perf_counter(); run(4, **kw); t4 = time.perf_counter() - t0
    t0 = time.perf_counter(); run(12, **kw); t12 = time.perf_counter() - t0
    print(f"{name:24s} steady-state cycle {(t12 - t4) / 8 * 1e3:7.2f} ms   (12 steps: {t12 / 12 * 1e3:.2f} ms/step)", flush=True)
