"""Synthetic insertion read sets for K3 (the POA consensus kernel): what the reads overlapping an INS record carry
between the insertion's flanks -- noisy copies of one inserted sequence.  Used by the tests, bench.py and
tools/make_poa_golden.py; seeded, numpy only.

Base codes are abPOA's (abPOA/src/abpoa_seq.c:15): A0 C1 G2 T3, anything else 4."""
from __future__ import annotations

import numpy as np

CODE = np.full(256, 4, np.uint8)
for _i, _c in enumerate("ACGT"):
    CODE[ord(_c)] = _i
    CODE[ord(_c.lower())] = _i


def encode(s: str) -> np.ndarray:
    return CODE[np.frombuffer(s.encode(), np.uint8)]


def decode(a) -> str:
    return "".join("ACGTN"[int(x)] for x in a)


def mutate(rng, s, err):
    """err/3 each of deletion, insertion, substitution per base (vectorised); never returns an empty read."""
    n = len(s)
    r = rng.random(n)
    keep = r >= err / 3
    sub = (r >= 2 * err / 3) & (r < err)
    ins = (r >= err / 3) & (r < 2 * err / 3)
    out = s.copy()
    out[sub] = (out[sub] + rng.integers(1, 4, int(sub.sum()))) % 4
    cnt = keep.astype(np.int64) + ins.astype(np.int64)
    idx = np.repeat(np.arange(n), cnt)
    res = out[idx]
    first = np.r_[True, idx[1:] != idx[:-1]] if len(idx) else np.zeros(0, bool)
    rnd = first & ins[idx]
    res[rnd] = rng.integers(0, 4, int(rnd.sum()))
    if len(res) == 0:
        res = np.array([0], np.uint8)
    return res.astype(np.uint8)


def read_set(rng, kind, length=None, n_reads=None, err=None):
    """One read set.  kind: 'random' (noisy copies of a random template), 'tandem' (tandem repeat with variable copy
    number -- the band-sensitive case), 'ragged' (truncated / extended / N-containing copies)."""
    if kind == "random":
        L = int(length if length is not None else rng.integers(50, 1500))
        tpl = rng.integers(0, 4, L).astype(np.uint8)
        e = err if err is not None else float(rng.choice([0.01, 0.05, 0.1, 0.15]))
        n = int(n_reads if n_reads is not None else rng.integers(2, 40))
        return [mutate(rng, tpl, e) for _ in range(n)]
    if kind == "tandem":
        unit = rng.integers(0, 4, int(rng.integers(2, 41))).astype(np.uint8)
        fl, fr = rng.integers(0, 4, 30).astype(np.uint8), rng.integers(0, 4, 30).astype(np.uint8)
        e = err if err is not None else float(rng.choice([0.03, 0.08, 0.15]))
        n = int(n_reads if n_reads is not None else rng.integers(8, 49))
        return [mutate(rng, np.concatenate([fl, np.tile(unit, int(rng.integers(20, 61))), fr]), e) for _ in range(n)]
    if kind == "ragged":
        L = int(length if length is not None else rng.integers(50, 600))
        tpl = rng.integers(0, 4, L).astype(np.uint8)
        n = int(n_reads if n_reads is not None else rng.integers(3, 30))
        out = []
        for _ in range(n):
            s = mutate(rng, tpl, err if err is not None else 0.08)
            if rng.random() < 0.3:
                s = s[:max(1, int(len(s) * rng.uniform(0.3, 1)))]
            if rng.random() < 0.2:
                s = np.concatenate([s, rng.integers(0, 5, int(rng.integers(1, 80))).astype(np.uint8)])
            out.append(s)
        return out
    raise ValueError(kind)


def pack(sets):
    """list of read sets (each a list of code arrays) -> (set_off, seq_off, bases) of svtk_poa_batch."""
    set_off = np.zeros(len(sets) + 1, np.int64)
    lens = []
    for i, s in enumerate(sets):
        set_off[i + 1] = set_off[i] + len(s)
        lens.extend(len(x) for x in s)
    seq_off = np.zeros(len(lens) + 1, np.int64)
    seq_off[1:] = np.cumsum(lens)
    flat = [x for s in sets for x in s]
    bases = np.concatenate(flat).astype(np.uint8) if flat else np.zeros(0, np.uint8)
    return set_off, seq_off, bases


def ins_workload(seed, n_sets, depth=60, carrier=0.8, err=0.08, len_lo=50, len_hi=10000, flank=0):
    """BASELINE.json configs[3]: INS records with inserted length log-uniform in [len_lo, len_hi], `depth`x coverage of
    which `carrier` of the reads carry the insertion, inserted copies mutated at `err` (ONT-like)."""
    rng = np.random.default_rng(seed)
    sets = []
    for _ in range(n_sets):
        L = int(round(np.exp(rng.uniform(np.log(len_lo), np.log(len_hi)))))
        tpl = rng.integers(0, 4, L + 2 * flank).astype(np.uint8)
        n = max(2, int(rng.binomial(depth, carrier)))
        sets.append([mutate(rng, tpl, err) for _ in range(n)])
    return sets


def ins_workload_shard(seed, n_total, world, rank, depth=60, carrier=0.8, err=0.08, len_lo=50, len_hi=10000):
    """Rank `rank`'s share of ONE job of n_total INS read sets (same distribution as ins_workload): every set's inserted length,
    read count and seed are drawn up front, the sets are dealt to the ranks heaviest first in snake order so that every rank gets
    the same estimated work (rows x band: reads x L x (20 + 0.02 L)), and only the rank's own sets are generated."""
    rng = np.random.default_rng(seed)
    L = np.rint(np.exp(rng.uniform(np.log(len_lo), np.log(len_hi), n_total))).astype(np.int64)
    n = np.maximum(2, rng.binomial(depth, carrier, n_total))
    seeds = rng.integers(0, 2**62, n_total)
    work = n * L * (20.0 + 0.02 * L)
    order = np.argsort(-work, kind="stable")
    k = np.arange(n_total)
    rnd, posn = k // world, k % world
    owner = np.where(rnd % 2 == 0, posn, world - 1 - posn)
    mine = order[owner == rank]
    sets = []
    for i in mine.tolist():
        r = np.random.default_rng(int(seeds[i]))
        tpl = r.integers(0, 4, int(L[i])).astype(np.uint8)
        sets.append([mutate(r, tpl, err) for _ in range(int(n[i]))])
    return sets, float(work[mine].sum()), float(work.sum())
