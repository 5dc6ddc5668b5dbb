"""NumPy transcription of k_finish's index arithmetic (segmented flip/disperse bitonic network with a
sub-range table), checked against the CPU builder.  A design aid for a box without a GPU: it validates
the algorithm, not the CUDA code."""
import sys
import numpy as np
import torch
sys.path.insert(0, ".")
import parallel_kd_tree_b200 as kd


def ordered(f):
    u = f.view(np.uint32).copy()
    u[(u << 1) == 0] = 0
    neg = (u & 0x80000000) != 0
    return np.where(neg, ~u, u | 0x80000000).astype(np.uint32)


def finish(pts, level=0):
    n, dim = pts.shape
    sidx = np.arange(n)
    sord = np.arange(n)
    stab = [(0, n)]
    j = 0
    while (n >> j) >= 3:
        nsub = 1 << j
        axis = (level + j) % dim
        skey = ordered(pts[sord, axis].copy())
        maxsort = (n >> j) - 1
        P, logHalf = 2, 0
        while P < maxsort:
            P <<= 1
            logHalf += 1
        halfP = P >> 1
        total = nsub << logHalf
        t = np.arange(total)
        sub = t >> logHalf
        w = t & (halfP - 1)
        tab = np.array(stab)
        slo, sn = tab[sub, 0], tab[sub, 1]
        k = 2
        while k <= P:
            jj = k >> 1
            while jj >= 1:
                flip = jj == (k >> 1)
                g, r = w // jj, w % jj
                if flip:
                    i = g * k + r
                    l = g * k + (k - 1 - r)
                else:
                    i = g * 2 * jj + r
                    l = i + jj
                ok = (sn >= 3) & (l < sn - 1)
                a, b = (slo + i)[ok], (slo + l)[ok]
                ka, kb = skey[a], skey[b]
                sw = (ka > kb) | ((ka == kb) & (sidx[sord[a]] > sidx[sord[b]]))
                a, b = a[sw], b[sw]
                skey[a], skey[b] = skey[b].copy(), skey[a].copy()
                sord[a], sord[b] = sord[b].copy(), sord[a].copy()
                jj >>= 1
            k <<= 1
        new = []
        for (lo, sn_) in stab:
            m = sn_ >> 1
            new.append((lo, m))
            new.append((lo + m + 1, sn_ - m - 1) if sn_ else (lo, 0))
        stab = new
        j += 1
    return sidx[sord]


for n, dim, seed in [(1, 2, 1), (2, 2, 1), (3, 3, 1), (4, 3, 2), (7, 2, 3), (100, 3, 4), (1000, 3, 5), (4095, 3, 6),
                     (4096, 2, 7), (3000, 1, 8), (2049, 4, 9)]:
    pts, _ = kd.ops.generate_problem(seed, dim, n, 0)
    got = finish(pts.numpy())
    ref = kd.ops.build(pts)[0].numpy()
    print(n, dim, "OK" if np.array_equal(got, ref) else "MISMATCH")
g = torch.Generator().manual_seed(0)
pts = torch.randint(-2, 3, (4096, 3), generator=g).float()
print("ties", "OK" if np.array_equal(finish(pts.numpy()), kd.ops.build(pts)[0].numpy()) else "MISMATCH")
