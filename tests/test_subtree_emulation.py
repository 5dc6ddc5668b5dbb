"""The arithmetic of k_subtree's team select (nibble counters, byte-prefix bin search, range-aligned digit schedule),
transcribed to NumPy in tools/emulate_subtree_select.py, against a plain sort: runs without a GPU."""
import sys
from pathlib import Path

import numpy as np
import pytest

sys.path.insert(0, str(Path(__file__).resolve().parent.parent / "tools"))
import emulate_subtree_select as em  # noqa: E402


def _want(f, idx, rank):
    key = em.ordered_key(f)
    return sorted((int(k) << 32) | int(i) for k, i in zip(key, idx))[rank]


@pytest.mark.parametrize("n", [9, 16, 33, 100, 200, 255])
def test_select_equals_sort_on_random_and_tied_keys(n):
    rng = np.random.default_rng(n)
    cases = [rng.uniform(-100, 100, n), rng.integers(-2, 3, n).astype(float), np.zeros(n),
             rng.choice([0.0, -0.0, 1e-30, -1e-30, 3e38, -3e38, 1.0], n), rng.uniform(50, 50.001, n)]
    for f in cases:
        f = f.astype(np.float32)
        idx = rng.permutation(1 << 20)[:n]
        for rank in (0, n // 2, n - 1):
            got, passes = em.team_select(f, idx, rank)
            assert got == _want(f, idx, rank)
            assert passes <= 16          # 8 key digits + 8 index digits at most


def test_bin_search_matches_a_cumulative_scan():
    rng = np.random.default_rng(1)
    for _ in range(200):
        counts = rng.multinomial(int(rng.integers(1, 256)), np.ones(16) / 16)
        digits = np.repeat(np.arange(16), counts)
        rng.shuffle(digits)
        accs = [em.thread_counts(list(digits[i:i + 8])) for i in range(0, len(digits), 8)]
        w = em.team_words(accs)
        cum = np.concatenate([[0], np.cumsum(counts)])
        for r in range(len(digits)):
            b, below, cnt = em.find_bin(w, r)
            want = int(np.searchsorted(cum, r, side="right") - 1)
            assert (b, below, cnt) == (want, int(cum[want]), int(counts[want]))
        assert em.find_bin(w, len(digits)) is None
