"""NumPy transcription of k_medium's select: digits aligned below the bits shared by the range's min and max key."""
import numpy as np


def select(keys, idx, DB):
    n1 = len(keys); n = n1 + 1
    comp = (keys.astype(np.uint64) << np.uint64(32)) | idx.astype(np.uint64)
    kmin, kmax = int(keys.min()), int(keys.max())
    vary = kmin ^ kmax
    top = vary.bit_length()
    prefix = 0 if top >= 32 else ((kmin >> top) << top) << 32
    rank, hi, shift, done = n >> 1, 32 + top, 64, False
    passes = 0
    for _ in range(8):
        if done:
            break
        dsh = max(hi - DB, 32) if hi > 32 else max(hi - DB, 0)
        mask = (1 << (hi - dsh)) - 1
        if dsh >= 32:
            valid = np.ones(n1, bool) if hi >= 64 else (keys >> (hi - 32)) == (prefix >> hi)
            digit = (keys >> (dsh - 32)) & mask
        else:
            valid = keys == (prefix >> 32)
            if hi < 32:
                valid &= (idx >> hi) == ((prefix & 0xffffffff) >> hi)
            digit = (idx >> dsh) & mask
        h = np.bincount(digit[valid].astype(np.int64), minlength=1 << DB)
        c = np.cumsum(h) - h
        b = np.nonzero((rank >= c) & (rank < c + h))[0]
        assert len(b) == 1
        b = int(b[0])
        prefix |= b << dsh
        rank -= int(c[b]); shift = dsh
        done = h[b] == 1 or dsh == 0
        hi = dsh; passes += 1
    cs = comp >> np.uint64(shift); want = prefix >> shift
    order = np.argsort(comp, kind="stable")
    node = cs == want
    assert node.sum() == 1 and (cs < want).sum() == n >> 1
    assert np.nonzero(node)[0][0] == order[n >> 1]
    return passes


rng = np.random.default_rng(1)
for DB in (8, 11):
    for n1 in (2, 3, 600, 16383):
        for kind in ("rand", "narrow", "few", "const"):
            if kind == "rand":
                keys = rng.integers(0, 2**32, n1, dtype=np.uint64).astype(np.uint32)
            elif kind == "narrow":
                keys = (0xC2C80000 + rng.integers(0, 5000, n1)).astype(np.uint32)
            elif kind == "few":
                keys = (rng.integers(0, 5, n1) * 77777 + 0x80000000).astype(np.uint32)
            else:
                keys = np.full(n1, 0xC2C80000, np.uint32)
            idx = rng.permutation(n1 * 3)[:n1].astype(np.uint32)
            print(DB, n1, kind, "passes", select(keys, idx, DB))
