"""GPU parity of the streaming (signature-index) form of the path: k_index_reads + k_refine_indexed against the CPU
oracle, against the reference-shaped per-(window, read) walker, and on the edge cases that are specific to it
(index pool re-dimensioning, multi-segment chunks, deferred leading-clip walks, empty CIGARs, packed-layout check)."""
import numpy as np
import pytest

from oracle import binding as B
from svtrek_b200 import api
from svtrek_b200 import simreads as S

pytestmark = pytest.mark.gpu


def _readset(reads, L=3_000_000):
    """reads: list of (pos, [cigar words]) on one contig -> ReadSet (sorted by pos)."""
    reads = sorted(reads, key=lambda r: r[0])
    pos = np.array([r[0] for r in reads], dtype=np.uint32)
    ncig = np.array([len(r[1]) for r in reads], dtype=np.uint32)
    off = np.zeros(len(reads) + 1, dtype=np.uint64); off[1:] = np.cumsum(ncig)
    cigar = np.array([w for r in reads for w in r[1]], dtype=np.uint32)
    rs = S.ReadSet(np.array([0, len(reads)], dtype=np.int64), pos, None, ncig, off, cigar, [L])
    rs.end = S.endpos_from_cigar(pos, off, cigar)
    return rs


def _both_paths(auditor, w):
    """(indexed result, walker result, index stats) for a dict of window arrays, through device buffers."""
    import torch
    dev = torch.device("cuda", 0)
    t = {k: torch.from_numpy(np.ascontiguousarray(w[k]).view(np.int32 if w[k].dtype.itemsize == 4 else np.uint8)).to(dev)
         for k in ("chrom", "start", "end", "pos", "kind")}
    n = int(w["chrom"].shape[0])
    res = torch.empty(n, dtype=torch.int32, device=dev)
    args = (n, t["chrom"].data_ptr(), t["start"].data_ptr(), t["end"].data_ptr(), t["pos"].data_ptr(), t["kind"].data_ptr(), res.data_ptr())
    torch.cuda.synchronize()
    st = auditor.measure_index(*args)
    indexed = res.cpu().numpy().copy()
    wk = auditor.measure_windows(*args)
    walker = res.cpu().numpy().copy()
    return indexed, walker, st, wk


def _win(chrom, start, end, pos, kind):
    m = 0xFFFFFFFF
    return dict(chrom=np.asarray(chrom, dtype=np.int32), start=(np.asarray(start, dtype=np.int64) & m).astype(np.uint32),
                end=(np.asarray(end, dtype=np.int64) & m).astype(np.uint32), pos=(np.asarray(pos, dtype=np.int64) & m).astype(np.uint32),
                kind=np.asarray(kind, dtype=np.uint8))


def test_indexed_equals_walker_equals_oracle_c1(auditor):
    rs, truth = S.make_config("c1")
    auditor.load_reads(rs)
    auditor.set_params(**api.DEFAULT_PARAMS)
    w = S.windows_from_truth(truth)
    indexed, walker, st, wk = _both_paths(auditor, w)
    want = B.refine_batch(B.OracleReads(rs), w["chrom"], w["start"], w["end"], w["pos"], w["kind"])
    assert np.array_equal(indexed, want)
    assert np.array_equal(walker, want)
    assert st["n_signatures"] == wk["n_signatures"]          # both paths collect the same multiset sizes
    assert st["n_words"] == rs.cigar.shape[0] and st["n_reads"] == rs.n_reads
    assert st["index_bytes"] >= 4 * st["n_words"]


def test_index_regrow_when_late_chunks_find_their_pools_already_empty(auditor):
    """many more signature-dense chunks than fit the GPU at once (16 sub-pools x a few resident chunks), so chunks START
    after their pool ran dry: all their flushes fail, their last events sit in their own slots uncounted, and the exact-size rebuild
    of the first retry can fall short.  The re-dimensioning now retries with slack until the build goes through; results == oracle."""
    rng = np.random.default_rng(8)
    reads = []
    for i in range(24_000):                                   # 3 000 chunks of 8 reads, ~40 events each: 3x the first-guess pool
        pos = int(100_000 + i * 40)
        words = []
        for _k in range(40):
            words.append(int(rng.integers(5, 30)) << 4)
            words.append(int(rng.integers(51, 60)) << 4 | (2 if rng.random() < 0.6 else 1))
        words.append(50 << 4)
        reads.append((pos, words))
    rs = _readset(reads)
    auditor.load_reads(rs)
    auditor.set_params(wider_interval=3000, median_interval=1500, narrow_interval=400)
    centers = np.array([150_000, 400_000, 700_000, 1_000_000, 300_000, 900_000, 500_000])
    kinds = np.array([0, 0, 0, 1, 1, 2, 2])
    w = _win(np.ones(7), centers - 3000, centers + 500, centers - 7, kinds)
    indexed, walker, st, _ = _both_paths(auditor, w)
    want = B.refine_batch(B.OracleReads(rs), w["chrom"], w["start"], w["end"], w["pos"], w["kind"])
    assert list(indexed) == list(want) and list(walker) == list(want)
    assert st["n_events"] == 24_000 * 40
    auditor.set_params(**api.DEFAULT_PARAMS)


def test_index_pool_regrow_and_multi_segment_chunks(auditor):
    """Reads that are almost all D > 50 ops: far more events than the first-guess pool, chunks that flush many
    512-event segments, tiles in which every second word is a candidate."""
    rng = np.random.default_rng(5)
    reads = []
    for _ in range(300):
        pos = int(500_000 + rng.integers(0, 3000))
        words = []
        for _k in range(400):
            words.append(int(rng.integers(5, 30)) << 4)
            words.append(int(rng.integers(51, 70)) << 4 | (2 if rng.random() < 0.7 else 1))
        words.append(100 << 4)
        reads.append((pos, words))
    for i in range(200):                                      # ordinary reads around them
        reads.append((int(400_000 + i * 1500), [15000 << 4]))
    rs = _readset(reads)
    auditor.load_reads(rs)
    auditor.set_params(**api.DEFAULT_PARAMS)
    centers = np.array([505_000, 510_000, 520_000, 505_000, 512_345, 530_000, 505_000])
    kinds = np.array([0, 0, 0, 1, 1, 1, 2])
    w = _win(np.ones(7), centers - 20000, centers + 2000, centers - 50, kinds)
    indexed, walker, st, _ = _both_paths(auditor, w)
    want = B.refine_batch(B.OracleReads(rs), w["chrom"], w["start"], w["end"], w["pos"], w["kind"])
    assert list(indexed) == list(want) and list(walker) == list(want)
    assert st["n_events"] == 300 * 400
    assert st["n_segments"] > st["n_chunks"]                  # chunks flushed more than one segment
    # the host-buffer entry point takes the same re-dimensioning route on a fresh context
    with api.Auditor(0) as fresh:
        fresh.load_reads(rs)
        got = fresh.refine_windows(w["chrom"], w["start"], w["end"], w["pos"], w["kind"])
        assert list(got) == list(want)


def test_deferred_leading_clip_walks(auditor):
    """DEL-END windows holding leading-soft-clip reads whose walk stops inside the window (refinement.c:210-220 records the
    post-walk cursor): few (defer list) and many (> 32: rescan path)."""
    rng = np.random.default_rng(11)
    reads = []
    for n_clip, center in ((5, 200_000), (80, 600_000)):
        for _ in range(n_clip):
            pos = int(center + rng.integers(-1500, 1500))
            words = [int(rng.integers(100, 900)) << 4 | 4]
            for _k in range(40):
                words.append(int(rng.integers(50, 400)) << 4)
                words.append(int(rng.integers(1, 3)) << 4 | int(rng.choice([1, 2])))
            words.append(3000 << 4)
            reads.append((pos, words))
        for _ in range(30):
            reads.append((int(center - 9000 + rng.integers(0, 6000)), [15000 << 4]))
    rs = _readset(reads)
    auditor.load_reads(rs)
    auditor.set_params(consensus_min_count=1, consensus_interval_range=5000)
    centers = np.array([200_000, 600_000, 200_000, 600_000])
    w = _win(np.ones(4), centers - 2000, centers + 2000, centers, [1, 1, 0, 0])
    indexed, walker, _, _ = _both_paths(auditor, w)
    want = B.refine_batch(B.OracleReads(rs), w["chrom"], w["start"], w["end"], w["pos"], w["kind"], 1, 5, 5000)
    assert list(indexed) == list(want) and list(walker) == list(want)
    assert want[0] != -1 and want[1] != -1
    for i in range(2):
        _, osig = B.refine(B.OracleReads(rs), 1, 1, int(w["start"][i]), int(w["end"][i]), int(w["pos"][i]), 1, 5, 5000, want_sig=True)
        assert len(osig) >= (5, 80)[i]
    auditor.set_params(**api.DEFAULT_PARAMS)


def test_empty_cigars_and_chunk_boundaries(auditor):
    """Reads with n_cigar == 0 (no signature), chunk sizes around the 8- and 32-read boundaries, single-read sets."""
    for n_reads in (1, 7, 8, 9, 31, 32, 33, 65):
        reads = []
        for i in range(n_reads):
            pos = 100_000 + 37 * i
            if i % 5 == 3:
                reads.append((pos, []))
            else:
                reads.append((pos, [200 << 4, 80 << 4 | 2, 300 << 4, 60 << 4 | 1, 400 << 4, 50 << 4 | 4]))
        rs = _readset(reads, L=1_000_000)
        rs.end = np.where(rs.n_cigar == 0, rs.pos + 1, rs.end).astype(np.uint32)
        auditor.load_reads(rs)
        auditor.set_params(consensus_min_count=1)
        w = _win(np.ones(3), [90_000, 99_000, 95_000], [110_000, 103_000, 105_000], [100_300, 100_500, 100_600], [0, 1, 2])
        indexed, walker, st, _ = _both_paths(auditor, w)
        want = B.refine_batch(B.OracleReads(rs), w["chrom"], w["start"], w["end"], w["pos"], w["kind"], 1, 5, 500)
        assert list(indexed) == list(want) and list(walker) == list(want), n_reads
        assert st["n_chunks"] == (n_reads + 7) // 8
    auditor.set_params(**api.DEFAULT_PARAMS)


def test_unpacked_layout_is_rejected(auditor):
    rs, _ = S.simulate(3, [200_000], 10, 8000, 0.005, 5, margin=30000)
    bad = S.ReadSet(rs.tid_off, rs.pos, rs.end, rs.n_cigar, rs.cigar_off.copy(), rs.cigar, rs.contig_len)
    bad.cigar_off[5] += 1
    with api.Auditor(0) as a:
        with pytest.raises(api.SvtkError, match="packed"):
            a.load_reads(bad)


def test_index_reused_across_batches_and_params(auditor):
    """The index depends on the reads only: parameter changes and later batches reuse it and stay exact."""
    rs, truth = S.simulate(21, [600_000], 40, 12000, 0.005, 150, margin=30000)
    auditor.load_reads(rs)
    orc = B.OracleReads(rs)
    for params in (dict(), dict(consensus_interval=1, consensus_min_count=2), dict(narrow_interval=500, wider_interval=5000)):
        p = dict(api.DEFAULT_PARAMS); p.update(params)
        auditor.set_params(**p)
        w = S.windows_from_truth(truth, wider=p["wider_interval"], median=p["median_interval"], narrow=p["narrow_interval"])
        got = auditor.refine_windows(w["chrom"], w["start"], w["end"], w["pos"], w["kind"])
        want = B.refine_batch(orc, w["chrom"], w["start"], w["end"], w["pos"], w["kind"], p["consensus_min_count"],
                              p["consensus_interval"], p["consensus_interval_range"])
        assert np.array_equal(got, want)
    auditor.set_params(**api.DEFAULT_PARAMS)
