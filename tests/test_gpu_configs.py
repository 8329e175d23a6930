"""GPU parity on BASELINE.json's synthetic configs (seeded) against the pinned CPU oracle, plus edge cases,
the overflow tiers of the signature segment, and size-independent properties at full config size."""
import numpy as np
import pytest

from oracle import binding as B
from svtrek_b200 import api
from svtrek_b200 import simreads as S

pytestmark = pytest.mark.gpu


def _check_config(auditor, rs, truth, params=None, sample=None):
    p = dict(api.DEFAULT_PARAMS); p.update(params or {})
    auditor.set_params(**p)
    w = S.windows_from_truth(truth, wider=p["wider_interval"], median=p["median_interval"], narrow=p["narrow_interval"])
    got = auditor.refine_windows(w["chrom"], w["start"], w["end"], w["pos"], w["kind"])
    idx = np.arange(len(got)) if sample is None else np.sort(np.random.default_rng(0).choice(len(got), sample, replace=False))
    orc = B.OracleReads(rs)
    want = B.refine_batch(orc, w["chrom"][idx], w["start"][idx], w["end"][idx], w["pos"][idx], w["kind"][idx],
                          p["consensus_min_count"], p["consensus_interval"], p["consensus_interval_range"])
    bad = np.flatnonzero(got[idx] != want)
    assert bad.size == 0, [(int(idx[b]), int(got[idx[b]]), int(want[b])) for b in bad[:8]]
    return got, w


def test_c1_all_windows(auditor):
    rs, truth = S.make_config("c1")
    auditor.load_reads(rs)
    got, _ = _check_config(auditor, rs, truth)
    assert (got != -1).mean() > 0.5          # the planted SVs are actually found
    for params in (dict(consensus_interval=1, consensus_min_count=2), dict(consensus_interval=20, consensus_min_count=10),
                   dict(consensus_interval_range=60), dict(wider_interval=3000, median_interval=1500, narrow_interval=400)):
        _check_config(auditor, rs, truth, params)
    auditor.set_params(**api.DEFAULT_PARAMS)


def test_c5_stress_sweep(auditor):
    """200x depth over tandem clusters: dense overlapping windows, thousands of signatures per window --
    exercises the overflow tiers; consensus-interval x min-count sweep of BASELINE.json configs[4]."""
    rs, truth = S.make_config("c5", scale=0.25)
    auditor.load_reads(rs)
    for interval in (1, 5, 20):
        for minc in (2, 3, 10):
            _check_config(auditor, rs, truth, dict(consensus_interval=interval, consensus_min_count=minc), sample=300)
    auditor.set_params(**api.DEFAULT_PARAMS)


def test_c4_ont_like_long_cigars(auditor):
    """ONT-like error profile: 1-3k CIGAR words per read => many 128-word warp steps per read."""
    rs, truth = S.make_config("c4", scale=0.02)
    assert rs.n_cigar.mean() > 1000
    auditor.load_reads(rs)
    _check_config(auditor, rs, truth)


def test_multi_contig(auditor):
    rs, truth = S.simulate(77, [400_000, 300_000, 250_000], 20, 12000, 0.005, 120, margin=30000)
    auditor.load_reads(rs)
    _check_config(auditor, rs, truth)


def test_edge_windows(auditor):
    rs, truth = S.make_config("c1")
    auditor.load_reads(rs)
    auditor.set_params(**api.DEFAULT_PARAMS)
    orc = B.OracleReads(rs)
    p0 = int(truth.pos0[10])
    cases = [  # chrom, start, end, pos, kind
        (1, 0, 5000, 100, 0), (1, 5000, 0, 100, 2), (0, p0 - 100, p0 + 100, p0, 0), (2, p0 - 100, p0 + 100, p0, 0), (-5, 1, 2, 1, 1),
        (1, p0 - 10000, p0 + 10000, p0, 3), (1, p0, p0, p0, 0), (1, p0, p0 + 1, p0, 2), (1, 4294967000, 4294967290, 5, 1),
        (1, 1, 4294967295, p0, 0), (1, 1, 999_999, p0, 2), (1, 900_000, 2_000_000, 950_000, 0), (1, p0 - 2, p0 + 2, p0, 0)]
    c = np.array(cases, dtype=np.int64)
    got = auditor.refine_windows(c[:, 0], c[:, 1], c[:, 2], c[:, 3], c[:, 4])
    want = B.refine_batch(orc, c[:, 0], c[:, 1] & 0xFFFFFFFF, c[:, 2] & 0xFFFFFFFF, c[:, 3] & 0xFFFFFFFF, c[:, 4])
    assert list(got) == list(want)
    # empty batch, and a batch before any reads exist
    assert len(auditor.refine_windows([], [], [], [], [])) == 0
    fresh = api.Auditor(0)
    with pytest.raises(api.SvtkError):
        fresh.refine_windows([1], [1], [100], [50], [0])
    fresh.close()


def test_overflow_tiers(auditor):
    """Windows with more signatures than the fast path's shared segment (1024), than the big shared segment
    (32768) and beyond (global arena)."""
    rng = np.random.default_rng(9)
    reads = []
    L = 3_000_000

    def add(n_reads, center, per_read):
        for _ in range(n_reads):
            pos = int(center - 9000 + rng.integers(0, 2000))
            words, ref = [], pos
            for _k in range(per_read):
                m = int(rng.integers(20, 60)); words.append(m << 4); ref += m
                words.append(int(rng.integers(51, 90)) << 4 | 2)
            words.append(500 << 4)
            reads.append((pos, words))
    add(40, 100_000, 3)          # ~120 signatures (fast path)
    add(300, 300_000, 12)        # ~3.6k (big shared segment)
    add(700, 600_000, 100)       # ~70k (global arena)
    add(400, 900_000, 100)       # ~40k: between 2^15 and the old 51200-int segment (bitonic padding to 2^16 must not fit in shared)
    reads.sort(key=lambda r: r[0])
    pos = np.array([r[0] for r in reads], dtype=np.uint32)
    ncig = np.array([len(r[1]) for r in reads], dtype=np.uint32)
    off = np.zeros(len(reads) + 1, dtype=np.uint64); off[1:] = np.cumsum(ncig)
    cigar = np.array([w for r in reads for w in r[1]], dtype=np.uint32)
    rs = S.ReadSet(np.array([0, len(reads)], dtype=np.int64), pos, None, ncig, off, cigar, [L])
    rs.end = S.endpos_from_cigar(pos, off, cigar)
    auditor.load_reads(rs)
    auditor.set_params(**api.DEFAULT_PARAMS)
    orc = B.OracleReads(rs)
    centers = np.array([100_000, 300_000, 600_000, 300_000, 600_000, 100_000, 900_000, 900_000])
    kinds = np.array([0, 0, 0, 1, 1, 1, 0, 1])
    st, en, ps = centers - 20000, centers + 20000, centers - 8000
    got = auditor.refine_windows(np.ones(8), st, en, ps, kinds)
    want = B.refine_batch(orc, np.ones(8), st, en, ps, kinds)
    assert list(got) == list(want)
    n_sig = [len(auditor.window_signatures(0, 1, int(s), int(e))) for s, e in zip(st[:3], en[:3])]
    assert n_sig[0] <= 1024 < n_sig[1] <= 32768 < n_sig[2], n_sig
    for i in range(3):
        _, osig = B.refine(orc, 0, 1, int(st[i]), int(en[i]), int(ps[i]), want_sig=True, cap=1 << 20)
        assert np.array_equal(auditor.window_signatures(0, 1, int(st[i]), int(en[i])), osig)


def test_c2_full_size_properties(auditor):
    """BASELINE.json configs[1] at full size (64 Mb, 128k reads, 10k records): idempotence, order independence,
    agreement of the batch seam with itself under window permutation, and the oracle on a 1500-window sample."""
    rs, truth = S.make_config("c2")
    auditor.load_reads(rs)
    got, w = _check_config(auditor, rs, truth, sample=1500)
    again = auditor.refine_windows(w["chrom"], w["start"], w["end"], w["pos"], w["kind"])
    assert np.array_equal(got, again)
    perm = np.random.default_rng(1).permutation(len(got))
    shuffled = auditor.refine_windows(w["chrom"][perm], w["start"][perm], w["end"][perm], w["pos"][perm], w["kind"][perm])
    assert np.array_equal(shuffled, got[perm])
    # every refined breakpoint lies within consensus-interval-range of the imprecise one (refinement.c:58,80)
    ok = got != -1
    assert (np.abs(got[ok].astype(np.int64) - w["pos"][ok].astype(np.int64)) < 500 + 5).all()
    assert ok.mean() > 0.5
