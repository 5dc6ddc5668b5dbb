"""Debugging aid for k_subtree: run the build with the shared-memory builder stopped after L sub-levels
(KD_B200_SUBTREE_MAXLEVELS, one fresh process per L) and report the first sub-level whose output is not
consistent with the CPU tree: node slots of finished levels must match exactly, every unfinished sub-range
must hold the right SET of points with the right pinned last element.

    python tools/debug_subtree.py N DIM [SEED]
"""
import os
import subprocess
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def child(n, dim, seed, L):
    import parallel_kd_tree_b200 as kd
    pts, _ = kd.ops.generate_problem(seed, dim, n, 0)
    ref = kd.ops.build(pts)[0]
    perm, _ = kd.ops.build(pts.cuda())
    torch.cuda.synchronize()
    got = perm.cpu()
    ranges = [(0, n)]
    for j in range(L):
        nxt = []
        for lo, sn in ranges:
            if sn == 0:
                continue
            m = sn // 2
            if sn >= 3 and got[lo + m] != ref[lo + m]:
                print(f"L={L}: level {j} range ({lo},{sn}): node slot {lo+m} got {int(got[lo+m])} want {int(ref[lo+m])}")
                return 1
            nxt += [(lo, m), (lo + m + 1, sn - m - 1)]
        ranges = nxt
    for lo, sn in ranges:
        if sn == 0:
            continue
        a, b = got[lo:lo + sn], ref[lo:lo + sn]
        if sorted(a.tolist()) != sorted(b.tolist()):
            print(f"L={L}: range ({lo},{sn}) holds the wrong set: {len(set(a.tolist()) ^ set(b.tolist()))} differ")
            return 1
        # the pinned last element of a range: the left-maximum rule only fixes it for ranges of >= 2 points
        want_last = None
    print(f"L={L}: ok ({len(ranges)} ranges)")
    return 0


if __name__ == "__main__":
    if len(sys.argv) > 4:
        sys.exit(child(int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])))
    n, dim = int(sys.argv[1]), int(sys.argv[2])
    seed = int(sys.argv[3]) if len(sys.argv) > 3 else n + dim
    for L in range(1, 14):
        env = dict(os.environ, KD_B200_SUBTREE_MAXLEVELS=str(L))
        r = subprocess.run([sys.executable, __file__, str(n), str(dim), str(seed), str(L)], env=env)
        if r.returncode:
            break
