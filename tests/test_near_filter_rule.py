"""The rule k_refine_warp uses to sort fewer signatures (svtk_index.cu): values further than range + interval + 1 from pos are
replaced by ONE representative per side (the largest below, the smallest above).  Property checked here on CPU with the oracle:
consensus_pos of the reduced array == consensus_pos of the full array, on the reference-derived golden cases and on random ones
(clusters straddling the reach boundary, empty sides, values at the very edge)."""
import os

import numpy as np

from oracle import binding as B
from tests import util


def reduce_like_the_kernel(v, pos, interval, rng):
    reach = rng + interval + 1
    if not (rng >= 0 and interval >= 0 and reach < (1 << 30) and 0 <= pos <= 0x7fffffff - reach - 1):
        return v
    v = np.asarray(v, dtype=np.int64)
    lo, hi = v[v < pos - reach], v[v > pos + reach]
    keep = v[(v >= pos - reach) & (v <= pos + reach)]
    return np.concatenate([[lo.max()] if lo.size else [], keep, [hi.min()] if hi.size else []]).astype(np.int32)


def test_golden_cases_survive_the_reduction():
    z = np.load(os.path.join(util.GOLD, "consensus_pos.npz"))
    vals, off, meta, exp = z["values"], z["offsets"], z["meta"], z["expected"]
    n = 0
    for i in range(len(exp)):
        pos, minc, interval, rng = (int(x) for x in meta[i])
        v = vals[off[i]:off[i + 1]]
        if len(v) < minc:
            continue
        red = reduce_like_the_kernel(v, pos, interval, rng)
        if len(red) >= minc:
            assert B.consensus_pos(red, pos, minc, interval, rng) == int(exp[i]), i
            n += 1
        else:
            assert int(exp[i]) == -1, i
    assert n > 1000


def test_random_cases_around_the_reach_boundary():
    rng = np.random.default_rng(5)
    n = 0
    for _ in range(4000):
        pos = int(rng.integers(2000, 1 << 30))
        interval = int(rng.choice([0, 1, 5, 20]))
        r = int(rng.choice([1, 30, 500, 3000]))
        minc = int(rng.choice([1, 2, 3, 5]))
        reach = r + interval + 1
        parts = []
        for _k in range(int(rng.integers(0, 5))):            # clusters anywhere, often right at the boundary
            c = pos + int(rng.choice([-reach - 2, -reach, -reach + 1, -r, -26, -25, 0, 25, 26, r - 1, r, reach, reach + 1, int(rng.integers(-3 * reach, 3 * reach))]))
            parts.append(c + rng.integers(-interval - 2, interval + 3, int(rng.integers(1, 8))))
        if rng.random() < 0.5:
            parts.append(rng.integers(0, 1 << 30, int(rng.integers(1, 20))))
        v = np.concatenate(parts).astype(np.int32) if parts else np.zeros(0, np.int32)
        red = reduce_like_the_kernel(v, pos, interval, r)
        if len(v) >= minc and len(red) >= minc:
            assert B.consensus_pos(red, pos, minc, interval, r) == B.consensus_pos(v, pos, minc, interval, r), (list(v), pos, minc, interval, r)
            n += 1
        elif len(v) >= minc:
            # fewer values left than min_count: no cluster can reach it (the kernel's scans find no candidate) -- and indeed
            assert B.consensus_pos(v, pos, minc, interval, r) == -1, (list(v), pos, minc, interval, r)
    assert n > 1500
