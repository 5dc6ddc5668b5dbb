"""NumPy transcription of k_finish v2: register/shuffle/shared bitonic network in XOR-partner form
(direction from bit k of the index inside the sub-range, last merge ascending, +inf padding)."""
import sys
import numpy as np
import torch
sys.path.insert(0, ".")
import parallel_kd_tree_b200 as kd
from emulate_finish import ordered

T = 4096
INF_KEY, INF_SLOT = np.uint32(0xFFFFFFFF), 0xFFFF


def finish(pts, level=0):
    n, dim = pts.shape
    sidx = np.arange(n)
    sord = np.arange(n)
    stab = [(0, n)]
    j = 0
    v = np.arange(T)
    while (n >> j) >= 3:
        nsub = 1 << j
        axis = (level + j) % dim
        maxsort = (n >> j) - 1
        P, logP = 2, 1
        while P < maxsort:
            P <<= 1
            logP += 1
        sub = v >> logP
        i = v & (P - 1)
        valid = sub < nsub
        tab = np.array(stab + [(0, 0)] * (T - len(stab)))
        slo, sn = tab[np.minimum(sub, T - 1), 0], tab[np.minimum(sub, T - 1), 1]
        inr = valid & (sn >= 3) & (i < sn - 1)
        pos = np.where(inr, slo + i, 0)
        slot = np.where(inr, sord[pos], INF_SLOT)
        key = np.where(inr, ordered(pts[np.where(inr, slot, 0), axis].copy()), INF_KEY).astype(np.uint32)

        def less(ak, aslot, bk, bslot):
            tie = ak == bk
            ia = np.where(aslot == INF_SLOT, 1 << 40, sidx[np.where(aslot == INF_SLOT, 0, aslot)])
            ib = np.where(bslot == INF_SLOT, 1 << 40, sidx[np.where(bslot == INF_SLOT, 0, bslot)])
            return (ak < bk) | (tie & (ia < ib))
        k = 2
        while k <= P:
            s = k >> 1
            while s >= 1:
                partner = v ^ s
                lower = (v & s) == 0
                asc = (k == P) | ((v & k) == 0)
                keep_min = lower == asc
                ok, oslot = key[partner], slot[partner]
                take = np.where(keep_min, less(ok, oslot, key, slot), less(key, slot, ok, oslot))
                key, slot = np.where(take, ok, key), np.where(take, oslot, slot)
                s >>= 1
            k <<= 1
        wr = inr
        sord[pos[wr]] = slot[wr]
        new = []
        for (lo, sn_) in stab:
            m = sn_ >> 1
            new.append((lo, m))
            new.append((lo + m + 1, sn_ - m - 1) if sn_ else (lo, 0))
        stab = new
        j += 1
    return sidx[sord]


if __name__ == "__main__":
    for n, dim, seed in [(1, 2, 1), (2, 2, 1), (3, 3, 1), (4, 3, 2), (7, 2, 3), (100, 3, 4), (1000, 3, 5),
                         (4095, 3, 6), (4096, 2, 7), (3000, 1, 8), (2049, 4, 9), (2100, 3, 10)]:
        pts, _ = kd.ops.generate_problem(seed, dim, n, 0)
        got = finish(pts.numpy())
        ref = kd.ops.build(pts)[0].numpy()
        print(n, dim, "OK" if np.array_equal(got, ref) else "MISMATCH")
    g = torch.Generator().manual_seed(0)
    pts = torch.randint(-2, 3, (4096, 3), generator=g).float()
    print("ties", "OK" if np.array_equal(finish(pts.numpy()), kd.ops.build(pts)[0].numpy()) else "MISMATCH")
