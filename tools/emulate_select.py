"""NumPy transcription of k_hist/k_scan/k_partition's digit, prefix and compare arithmetic for one range."""
import numpy as np

SH = [53, 42, 32, 21, 10, 0]
BITS = [11, 11, 10, 11, 11, 10]


def select(keys, idx):
    n1 = len(keys)  # unpinned elements
    n = n1 + 1
    prefix, rank, shift, done = 0, n >> 1, 64, False
    comp = (keys.astype(np.uint64) << np.uint64(32)) | idx.astype(np.uint64)
    for d in range(6):
        if done:
            break
        dsh, mask = SH[d], (1 << BITS[d]) - 1
        if d < 3:
            ks = dsh - 32
            valid = np.ones(n1, bool)
            if d > 0:
                ps = SH[d - 1] - 32
                want = prefix >> SH[d - 1]
                valid = (keys >> ps) == want
            digit = (keys >> ks) & mask
        else:
            wantkey = prefix >> 32
            valid = keys == wantkey
            if d > 3:
                ps = SH[d - 1]
                wantidx = (prefix & 0xffffffff) >> ps
                valid &= (idx >> ps) == wantidx
            digit = (idx >> dsh) & mask
        h = np.bincount(digit[valid], minlength=2048)
        c = np.cumsum(h) - h
        b = np.nonzero((rank >= c) & (rank < c + h))[0]
        assert len(b) == 1, (d, b)
        b = int(b[0])
        prefix |= b << dsh
        rank -= int(c[b])
        shift = dsh
        done = h[b] == 1 or d == 5
    want = prefix >> shift
    cs = comp >> np.uint64(shift)
    left, node, right = cs < want, cs == want, cs > want
    order = np.argsort(comp, kind="stable")
    assert node.sum() == 1 and left.sum() == n >> 1, (node.sum(), left.sum(), n >> 1)
    assert np.nonzero(node)[0][0] == order[n >> 1]
    return d


rng = np.random.default_rng(0)
for n1 in [2, 3, 10, 1000, 100000]:
    for kind in ["rand", "few", "const"]:
        if kind == "rand":
            keys = rng.integers(0, 2**32, n1, dtype=np.uint64).astype(np.uint32)
        elif kind == "few":
            keys = rng.integers(0, 5, n1, dtype=np.uint64).astype(np.uint32) * 77777 + 0x80000000
        else:
            keys = np.full(n1, 0xC2C80000, np.uint32)
        idx = rng.permutation(n1 * 3)[:n1].astype(np.uint32)
        print(n1, kind, "passes", select(keys, idx) + 0)
