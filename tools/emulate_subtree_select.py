"""NumPy transcription of the arithmetic of k_subtree's team select (csrc/build_kernels.cu, sub_team_level):
16 four-bit counters per thread in one 64-bit word (digit d in the low (d < 8) or high nibble of byte d & 7), the two
nibble planes widened to byte fields, byte prefix sums by one multiply per word, the digit schedule aligned to the live
keys' range (key - min, then index digits among equal keys), and the all-pairs finish.  Checked against a sort by
tests/test_subtree_emulation.py; running this file prints a few cases."""
import numpy as np

M64 = (1 << 64) - 1


def ordered_key(f):
    u = np.asarray(f, dtype=np.float32).view(np.uint32).astype(np.uint64)
    u = np.where((u << np.uint64(1)) & np.uint64(0xFFFFFFFF) == 0, np.uint64(0), u)
    return np.where(u & np.uint64(0x80000000), (~u) & np.uint64(0xFFFFFFFF), u | np.uint64(0x80000000)).astype(np.uint64)


def thread_counts(digits):
    """One thread's (<= 8) digits -> the 64-bit accumulator of the kernel."""
    acc = 0
    for d in digits:
        d = int(d)
        acc = (acc + (1 << (((d & 7) << 3) | ((d >> 3) << 2)))) & M64
    return acc


def team_words(accs):
    """Per-thread accumulators -> the four summed byte-field words (bins 4k .. 4k+3 in word k)."""
    w = [0, 0, 0, 0]
    for acc in accs:
        pa, pb = acc & 0x0F0F0F0F0F0F0F0F, (acc >> 4) & 0x0F0F0F0F0F0F0F0F
        for k, v in enumerate((pa & 0xFFFFFFFF, pa >> 32, pb & 0xFFFFFFFF, pb >> 32)):
            w[k] += v          # byte fields never overflow: a team holds at most 255 elements
    return w


def find_bin(w, r):
    """(bin, elements below it, elements in it) of rank r; the kernel's multiply trick."""
    p = [(x * 0x01010101) & 0xFFFFFFFF for x in w]
    c = [0, p[0] >> 24, (p[0] >> 24) + (p[1] >> 24), (p[0] >> 24) + (p[1] >> 24) + (p[2] >> 24)]
    total = c[3] + (p[3] >> 24)
    if r >= total:
        return None
    kk = sum(1 for t in c[1:] if r >= t)
    rr = r - c[kk]
    b = [(p[kk] >> (8 * i)) & 0xFF for i in range(4)]
    bi = sum(1 for t in b[:3] if rr >= t)
    below = 0 if bi == 0 else b[bi - 1]
    return 4 * kk + bi, c[kk] + below, b[bi] - below


def team_select(keys_f32, idx, rank, cap=8):
    """Composite (key, idx) of the element of 0-based `rank`, by the kernel's select (<= 255 elements)."""
    key = ordered_key(keys_f32)
    comp = (key << np.uint64(32)) | np.asarray(idx, dtype=np.uint64)
    alive = np.ones(len(comp), dtype=bool)
    kmin, kmax = int(key.min()), int(key.max())
    mode, base, sh, r, acnt, passes = 0, kmin, 0, rank, len(comp), 0
    if kmin == kmax:
        mode, sh = 1, 28
    else:
        t = (kmax - kmin).bit_length() - 1
        sh = t - 3 if t > 3 else 0
    while acnt > cap:
        passes += 1
        assert passes <= 20
        u = (comp & np.uint64(0xFFFFFFFF)) if mode else ((comp >> np.uint64(32)) - np.uint64(base))
        dig = ((u >> np.uint64(sh)) & np.uint64(15)).astype(int)
        accs = [thread_counts([dig[i] for i in range(t0, min(t0 + 8, len(comp))) if alive[i]]) for t0 in range(0, len(comp), 8)]
        tb, cbase, ccnt = find_bin(team_words(accs), int(r))
        r, acnt = r - cbase, ccnt
        alive &= dig == tb
        if mode == 0:
            base += tb << sh
            if sh == 0:
                mode, sh = 1, 28
            else:
                sh = sh - 4 if sh > 4 else 0
        else:
            sh = sh - 4 if sh >= 4 else 0
    cand = comp[alive]
    ranks = [int((cand < x).sum()) for x in cand]     # all-pairs finish
    return int(cand[ranks.index(r)]), passes


if __name__ == "__main__":
    rng = np.random.default_rng(0)
    for n in (9, 31, 100, 255):
        f = rng.uniform(-100, 100, n).astype(np.float32)
        got, passes = team_select(f, np.arange(n), n // 2)
        key = ordered_key(f)
        want = sorted((int(k) << 32) | i for i, k in enumerate(key))[n // 2]
        print(n, "ok" if got == want else "MISMATCH", "digit passes:", passes)
