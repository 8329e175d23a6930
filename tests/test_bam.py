"""svtk_load_bam: BGZF inflate + BAM record extraction on the GPU.

The reference reads BAM through HTSlib (hts_open / sam_index_load / sam_itr_queryi, audit.c:270-272, refinement.c:109-117),
which inflates BGZF blocks with zlib (bgzf.c:746).  Inflate output is fully determined by the input, so the bar is bit-exact:

  not gpu : the DEFLATE decoder source (svtk_bgzf_core.cuh) compiled as plain C++ equals zlib on stored / fixed / dynamic
            blocks of several data kinds and rejects corrupt streams; the host-side BGZF / BAI parsers agree with the Python
            restatement in tests/util_bam.py
  gpu     : svtk_load_bam on BAM files written by the reference's own HTSlib (oracle/_ref) yields exactly the reads that were
            written, and refinement results equal the oracle's on the same windows
"""
import ctypes as C
import os
import subprocess
import tempfile
import zlib

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def hostsim():
    d = tempfile.mkdtemp(prefix="svtk_hostsim_")
    so = os.path.join(d, "libbgzf_hostsim.so")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-o", so, os.path.join(ROOT, "tests", "hostsim", "bgzf_hostsim.cpp")])
    L = C.CDLL(so)
    L.hostsim_inflate.argtypes = [C.c_void_p, C.c_longlong, C.c_longlong, C.c_longlong, C.c_void_p, C.c_uint]
    return L


def raw_deflate(data, level, strategy=zlib.Z_DEFAULT_STRATEGY):
    c = zlib.compressobj(level, zlib.DEFLATED, -15, 9, strategy)
    return c.compress(data) + c.flush()


def payloads(rng, n):
    for t in range(n):
        ln = int(rng.integers(0, 65280)) if t else 0
        kind = t % 6
        if kind == 0:
            yield rng.integers(0, 256, ln, dtype=np.uint8).tobytes()
        elif kind == 1:
            yield bytes(rng.choice(np.frombuffer(b"ACGT", np.uint8), ln))
        elif kind == 2:
            yield (b"chr1\t12345\tACGTTTGA" * (ln // 19 + 1))[:ln]
        elif kind == 3:
            yield bytes(rng.integers(33, 74, ln, dtype=np.uint8))
        elif kind == 4:
            yield bytes(ln)
        else:
            a = rng.integers(0, 256, ln, dtype=np.uint8)
            a[rng.random(ln) < 0.7] = 65
            yield a.tobytes()


def run_inflate(L, comp, out_len, pad=0):
    """decode at input byte offset `pad` into an output slice that starts `pad + 1` bytes past an aligned address; the bytes
    around the slice (the neighbouring blocks' output on the GPU) must stay untouched"""
    buf = np.frombuffer(b"\xAA" * pad + bytes(comp) + b"\x55" * 5, np.uint8).copy()
    lead = 8 + pad + 1
    out = np.full(lead + out_len + 24, 0xEE, np.uint8)
    rc = L.hostsim_inflate(buf.ctypes.data, len(buf), pad, pad + len(comp), out.ctypes.data + lead, out_len)
    assert (out[:lead] == 0xEE).all() and (out[lead + out_len:] == 0xEE).all(), "decoder wrote outside its slice"
    return rc, out[lead:lead + out_len].tobytes()


def test_decoder_equals_zlib(hostsim):
    rng = np.random.default_rng(20261104)
    modes = ((6, zlib.Z_DEFAULT_STRATEGY), (1, zlib.Z_DEFAULT_STRATEGY), (9, zlib.Z_DEFAULT_STRATEGY), (0, zlib.Z_DEFAULT_STRATEGY),
             (6, zlib.Z_FIXED), (6, zlib.Z_HUFFMAN_ONLY), (6, zlib.Z_RLE), (6, zlib.Z_FILTERED))
    for t, data in enumerate(payloads(rng, 60)):
        for level, strat in modes:
            rc, out = run_inflate(hostsim, raw_deflate(data, level, strat), len(data), pad=t % 8)
            assert rc == 0, (t, level, strat, rc)
            assert out == data, (t, level, strat)


def test_decoder_rejects_or_matches_zlib_on_corrupt_streams(hostsim):
    rng = np.random.default_rng(7)
    data = rng.integers(0, 64, 30000, dtype=np.uint8).tobytes()
    comp = raw_deflate(data, 6)
    rejected = 0
    for _ in range(200):
        c2 = bytearray(comp)
        c2[int(rng.integers(0, len(c2)))] ^= 1 << int(rng.integers(0, 8))
        rc, out = run_inflate(hostsim, c2, len(data))
        try:
            d = zlib.decompressobj(-15)
            ref = d.decompress(bytes(c2))
            ref_ok = d.eof and len(ref) == len(data)
        except zlib.error:
            ref, ref_ok = None, False
        if rc == 0:
            assert ref_ok and out == ref          # a flip zlib also accepts (raw DEFLATE carries no checksum): same bytes
        else:
            rejected += 1
    assert rejected >= 1
    # wrong ISIZE, truncated payload
    assert run_inflate(hostsim, comp, len(data) - 1)[0] != 0
    assert run_inflate(hostsim, comp, len(data) + 1)[0] != 0
    assert run_inflate(hostsim, comp[:len(comp) // 2], len(data))[0] != 0


def test_sliced_crc32_equals_zlib(hostsim):
    """the CRC the loader checks every block with: per-slice CRCs merged by the x^(8n) identity == zlib.crc32"""
    hostsim.hostsim_crc32.argtypes = [C.c_void_p, C.c_uint, C.c_uint]
    hostsim.hostsim_crc32.restype = C.c_uint
    rng = np.random.default_rng(0)
    for n in (0, 1, 2, 31, 32, 33, 100, 255, 4096, 12345, 65280, 65535, 65536):
        d = rng.integers(0, 256, max(n, 1), dtype=np.uint8)
        for slices in (1, 2, 7, 32):
            assert hostsim.hostsim_crc32(d.ctypes.data, n, slices) == zlib.crc32(d[:n].tobytes()) & 0xFFFFFFFF, (n, slices)


# ---- golden BAM written by the reference's HTSlib (tools/make_bam_golden.py)
GOLD_BAM = os.path.join(ROOT, "tests", "golden", "bam_small.bam")
GOLD_NPZ = os.path.join(ROOT, "tests", "golden", "bam_small.npz")
FIELDS = ("tid_off", "pos", "end", "n_cigar", "cigar_off", "cigar")


def test_python_reader_matches_golden():
    from tests import util_bam as U
    z = np.load(GOLD_NPZ)
    got = U.read_bam(GOLD_BAM)
    for k in FIELDS:
        assert np.array_equal(got[k], z[k]), k
    assert [n for n, _ in got["refs"]] == list(z["ref_names"])
    assert int(z["n_cigar"].max()) == 70001 and int((z["n_cigar"] == 0).sum()) >= 2      # CG-tag read, `*` CIGARs


def test_decoder_inflates_every_block_of_the_golden_bam(hostsim):
    from tests import util_bam as U
    buf = open(GOLD_BAM, "rb").read()
    blocks = U.bgzf_blocks(buf)
    assert len(blocks) > 10 and blocks[-1][3] == 0                                     # ends with the empty EOF block
    for off, p, n, isize in blocks:
        rc, out = run_inflate(hostsim, buf[p:p + n], isize, pad=off % 4)
        assert rc == 0 and out == zlib.decompress(buf[p:p + n], -15), off


# ---- GPU
def _same_reads(got, want):
    for k in FIELDS:
        assert np.array_equal(np.asarray(got[k]).astype(np.int64), np.asarray(want[k]).astype(np.int64)[:len(got[k])]), k


@pytest.mark.gpu
def test_load_bam_golden():
    from svtrek_b200 import api
    z = np.load(GOLD_NPZ)
    a = api.Auditor(0)
    st = a.load_bam(GOLD_BAM)
    got = a.copy_reads()
    for k in FIELDS:
        assert np.array_equal(got[k].astype(np.int64), z[k].astype(np.int64)), k
    assert a.bam_refs() == list(zip([str(x) for x in z["ref_names"]], [int(x) for x in z["ref_lens"]]))
    assert st["n_reads"] == len(z["pos"]) and st["n_words"] == len(z["cigar"]) and st["n_blocks"] == 19 and st["n_ref"] == 3
    assert st["inflated_bytes"] > st["file_bytes"]
    # loading again (buffers reused) gives the same thing
    a.load_bam(GOLD_BAM, GOLD_BAM + ".bai")
    got2 = a.copy_reads()
    for k in FIELDS:
        assert np.array_equal(got2[k], got[k]), k


@pytest.mark.gpu
def test_load_bam_errors(tmp_path):
    from svtrek_b200 import api
    a = api.Auditor(0)
    buf = bytearray(open(GOLD_BAM, "rb").read())
    bai = open(GOLD_BAM + ".bai", "rb").read()

    def attempt(name, data, index=bai):
        p = tmp_path / name
        p.write_bytes(bytes(data))
        if index is not None:
            (tmp_path / (name + ".bai")).write_bytes(index)
        with pytest.raises(api.SvtkError):
            a.load_bam(str(p))

    attempt("noindex.bam", buf, index=None)
    attempt("trunc.bam", buf[:len(buf) // 2])
    attempt("magic.bam", b"\0" + bytes(buf[1:]))
    from tests import util_bam as U
    off, p, n, isize = U.bgzf_blocks(bytes(buf))[3]
    bad = bytearray(buf)
    bad[p + 2:p + 40] = bytes(38)                      # wipe the start of a DEFLATE payload: dynamic-block header becomes garbage
    attempt("payload.bam", bad)
    import struct
    attempt("nref.bam", buf, index=bai[:4] + struct.pack("<i", 2) + bai[8:])                       # index of another file: reference count differs
    assert struct.unpack_from("<i", bai, 8)[0] > 0 and struct.unpack_from("<i", bai, 16)[0] > 0       # ref 0 has a bin with a chunk at byte 20
    attempt("stale1.bam", buf, index=bai[:20] + struct.pack("<Q", 12345 << 16) + bai[28:])          # chunk start inside no block
    attempt("stale2.bam", buf, index=bai[:20] + struct.pack("<Q", off << 16 | 7) + bai[28:])        # a real block, but not a record boundary
    # a flip that still inflates to the right length (raw DEFLATE has no checksum): only the block's CRC32 catches it
    rng = np.random.default_rng(5)
    off, p, n, isize = U.bgzf_blocks(bytes(buf))[5]
    good = zlib.decompress(bytes(buf[p:p + n]), -15)
    for _ in range(4000):
        c2 = bytearray(buf)
        c2[p + int(rng.integers(n // 4, n))] ^= 1 << int(rng.integers(0, 8))
        try:
            d = zlib.decompressobj(-15)
            alt = d.decompress(bytes(c2[p:p + n]))
            if d.eof and len(alt) == isize and alt != good and not d.unused_data:
                break
        except zlib.error:
            pass
    else:
        pytest.fail("no silent flip found")
    (tmp_path / "crc.bam").write_bytes(bytes(c2))
    (tmp_path / "crc.bam.bai").write_bytes(bai)
    with pytest.raises(api.SvtkError, match="CRC32"):
        a.load_bam(str(tmp_path / "crc.bam"))
    with pytest.raises(api.SvtkError):
        a.load_bam(str(tmp_path / "does_not_exist.bam"))
    # the ctx is still usable afterwards
    a.load_bam(GOLD_BAM)
    assert a.copy_reads()["pos"].shape[0] == 256


@pytest.mark.gpu
def test_load_bam_live_equals_written_reads_and_refines_like_the_oracle(tmp_path):
    """HTSlib-written BAM of a C1-like read set: loader output == the reads that were written; windows refined on the loaded
    reads == the oracle on the original arrays."""
    from oracle import binding as B
    from svtrek_b200 import api, simreads as S
    from tests import util_bam as U
    if not os.path.exists(os.path.join(B.REF_DIR, "test_view")):
        pytest.skip("oracle/_ref HTSlib tools not built here")
    rs, truth = S.make_config("c1", scale=0.25)
    sam, bam = str(tmp_path / "c1.sam"), str(tmp_path / "c1.bam")
    S.write_sam(sam, rs)
    B.make_bam(sam, bam)
    a = api.Auditor(0)
    st = a.load_bam(bam)
    got = a.copy_reads()
    assert st["n_reads"] == rs.n_reads and st["n_pieces"] >= 2
    want = dict(tid_off=rs.tid_off, pos=rs.pos, end=rs.end, n_cigar=rs.n_cigar, cigar_off=rs.cigar_off[:-1], cigar=rs.cigar)
    for k in FIELDS:
        assert np.array_equal(got[k].astype(np.int64), np.asarray(want[k]).astype(np.int64)), k
    py = U.read_bam(bam)
    for k in FIELDS:
        assert np.array_equal(got[k].astype(np.int64), py[k].astype(np.int64)), k
    w = S.windows_from_truth(truth)
    res = a.refine_windows(w["chrom"], w["start"], w["end"], w["pos"], w["kind"])
    want = B.refine_batch(B.OracleReads(rs), w["chrom"], w["start"], w["end"], w["pos"], w["kind"])
    assert np.array_equal(res, want) and int((want >= 0).sum()) > 10


def test_python_bam_writer_round_trips_and_htslib_reads_it(tmp_path):
    """simreads.write_bam (the bench's file-level input): the plain-Python reader gets the reads back; where the reference's
    HTSlib tools exist, they render the same POS / CIGAR from it."""
    from oracle import binding as B
    from svtrek_b200 import simreads as S
    from tests import util_bam as U
    rs, _ = S.make_config("c1", scale=0.25)
    bam = str(tmp_path / "w.bam")
    S.write_bam(bam, rs)
    got = U.read_bam(bam)
    want = dict(tid_off=rs.tid_off, pos=rs.pos, end=rs.end, n_cigar=rs.n_cigar, cigar_off=rs.cigar_off[:-1], cigar=rs.cigar)
    for k in FIELDS:
        assert np.array_equal(got[k].astype(np.int64), np.asarray(want[k]).astype(np.int64)), k
    if os.path.exists(os.path.join(B.REF_DIR, "test_view")):
        txt = subprocess.run([os.path.join(B.REF_DIR, "test_view"), bam], stdout=subprocess.PIPE, check=True).stdout.decode().splitlines()
        rows = [l.split("\t") for l in txt if not l.startswith("@")]
        assert len(rows) == rs.n_reads
        for r in (0, 1, len(rows) // 2, len(rows) - 1):
            assert int(rows[r][3]) - 1 == int(rs.pos[r])
            assert rows[r][5] == S.cigar_string(rs.cigar[int(rs.cigar_off[r]):int(rs.cigar_off[r + 1])])


@pytest.mark.gpu
def test_load_bam_from_python_writer_multi_contig(tmp_path):
    """three contigs (one without reads), loader == written reads; refinement on the loaded reads == oracle"""
    from oracle import binding as B
    from svtrek_b200 import api, simreads as S
    rs, truth = S.simulate(11, contig_lens=[400_000, 300_000], coverage=12, read_len=6000, small_rate=0.01, n_sv=60, min_gap=1500)
    # splice an empty contig in the middle: tid 0, (empty) 1, 2
    rs3 = S.ReadSet(tid_off=np.array([rs.tid_off[0], rs.tid_off[1], rs.tid_off[1], rs.tid_off[2]], np.int64), pos=rs.pos, end=rs.end,
                    n_cigar=rs.n_cigar, cigar_off=rs.cigar_off, cigar=rs.cigar, contig_len=[rs.contig_len[0], 50_000, rs.contig_len[1]])
    bam = str(tmp_path / "m.bam")
    S.write_bam(bam, rs3)
    a = api.Auditor(0)
    st = a.load_bam(bam)
    got = a.copy_reads()
    assert st["n_ref"] == 3 and a.bam_refs()[1] == ("chr2", 50_000)
    want = dict(tid_off=rs3.tid_off, pos=rs3.pos, end=rs3.end, n_cigar=rs3.n_cigar, cigar_off=rs3.cigar_off[:-1], cigar=rs3.cigar)
    for k in FIELDS:
        assert np.array_equal(got[k].astype(np.int64), np.asarray(want[k]).astype(np.int64)), k
    truth.tid[truth.tid == 1] = 2
    w = S.windows_from_truth(truth)
    res = a.refine_windows(w["chrom"], w["start"], w["end"], w["pos"], w["kind"])
    want_r = B.refine_batch(B.OracleReads(rs3), w["chrom"], w["start"], w["end"], w["pos"], w["kind"])
    assert np.array_equal(res, want_r) and int((want_r >= 0).sum()) > 5


@pytest.mark.gpu
def test_ins_sequences_equal_the_restatement_and_feed_the_poa(tmp_path):
    """svtk_ins_sequences: read sets == the plain-Python restatement on the same BAM; their POA consensus
    (svtk_poa_consensus) == abPOA's on the same sets and recovers the planted insertion."""
    from oracle import binding as B
    from svtrek_b200 import api, simreads as S
    from tests import util_bam as U
    rs, truth = S.make_config("c1", scale=0.25)
    bam = str(tmp_path / "i.bam")
    S.write_bam(bam, rs, truth=truth)
    a = api.Auditor(0)
    a.load_bam(bam)
    w = S.windows_from_truth(truth)
    refined = a.refine_windows(w["chrom"], w["start"], w["end"], w["pos"], w["kind"])
    got = a.ins_sequences(w["chrom"], w["start"], w["end"], w["pos"], w["kind"], refined, tol=5)
    rd = U.read_bam(bam)
    want = U.ins_sequences(rd, w["chrom"], w["start"], w["end"], w["kind"], refined, tol=5)
    assert len(want) > 20 and len(got["window"]) == len(want)
    for s, (win, seqs) in enumerate(want):
        lo, hi = int(got["set_off"][s]), int(got["set_off"][s + 1])
        assert int(got["window"][s]) == win and hi - lo == len(seqs)
        for k, (r, rp, codes) in enumerate(seqs):
            assert int(got["read"][lo + k]) == r and int(got["ref_pos"][lo + k]) == rp
            assert np.array_equal(got["bases"][int(got["seq_off"][lo + k]):int(got["seq_off"][lo + k + 1])], codes), (win, r)
    # straight into K3
    cons, st = a.poa_consensus(got["set_off"], got["seq_off"], got["bases"])
    assert st["n_sets"] == len(want)
    ins = np.flatnonzero(~truth.is_del)
    rec_of_win = w["record"]
    checked = 0
    for s, (win, seqs) in enumerate(want):
        if len(seqs) < 5:
            continue
        L = int(truth.length[int(rec_of_win[win])])
        assert abs(len(cons[s]) - L) <= max(2, L // 50), (win, len(cons[s]), L)
        if B.have_abpoa() and checked < 6:
            ref_cons, _ = B.ref_poa_msa([c for _, _, c in seqs])
            assert np.array_equal(cons[s], ref_cons), win
            checked += 1
    assert len(ins) > 0
    # needs the BAM image: decoded-array loads cannot serve it
    a.load_reads(rs)
    with pytest.raises(api.SvtkError):
        a.ins_sequences(w["chrom"], w["start"], w["end"], w["pos"], w["kind"], refined)


@pytest.mark.gpu
@pytest.mark.parametrize("case", ["empty", "ont", "contigs24", "deep"])
def test_load_bam_shapes(tmp_path, case):
    """header-only file; ONT-like reads (1-3 k CIGAR words each); 24 references, some without reads; a 200x pile-up (long record
    chains between two index offsets).  Loader == written reads, refinement == oracle."""
    from oracle import binding as B
    from svtrek_b200 import api, simreads as S
    if case == "empty":
        rs = S.ReadSet(tid_off=np.zeros(3, np.int64), pos=np.zeros(0, np.uint32), end=np.zeros(0, np.uint32), n_cigar=np.zeros(0, np.uint32),
                       cigar_off=np.zeros(1, np.uint64), cigar=np.zeros(0, np.uint32), contig_len=[1000, 2000])
        truth = None
    elif case == "ont":
        rs, truth = S.simulate(3, contig_lens=[300_000], coverage=20, read_len=15000, small_rate=0.08, n_sv=40, sv_kind="ins", min_len=50, max_len=3000,
                               log_uniform=True, min_gap=1500)
    elif case == "contigs24":
        rs, truth = S.simulate(4, contig_lens=[250_000 if t % 5 else 200_000 for t in range(24)], coverage=4, read_len=8000, small_rate=0.01, n_sv=120, min_gap=1500)
    else:
        rs, truth = S.simulate(5, contig_lens=[250_000], coverage=200, read_len=5000, small_rate=0.005, n_sv=30, min_gap=1500)
    bam = str(tmp_path / "x.bam")
    S.write_bam(bam, rs, truth=truth)
    a = api.Auditor(0)
    st = a.load_bam(bam)
    got = a.copy_reads()
    want = dict(tid_off=rs.tid_off, pos=rs.pos, end=rs.end, n_cigar=rs.n_cigar, cigar_off=rs.cigar_off[:-1], cigar=rs.cigar)
    for k in FIELDS:
        assert np.array_equal(got[k].astype(np.int64), np.asarray(want[k]).astype(np.int64)), (case, k)
    assert st["n_reads"] == rs.n_reads
    if truth is None:
        assert a.refine_windows([1], [100], [900], [500], [2])[0] == -1
        return
    if case == "ont":
        assert int(rs.n_cigar.max()) > 1000
    w = S.windows_from_truth(truth)
    res = a.refine_windows(w["chrom"], w["start"], w["end"], w["pos"], w["kind"])
    assert np.array_equal(res, B.refine_batch(B.OracleReads(rs), w["chrom"], w["start"], w["end"], w["pos"], w["kind"])), case


@pytest.mark.gpu
def test_load_bam_refs_shards_equal_the_whole(tmp_path):
    """svtk_load_bam_refs: every reference group's reads == that slice of the whole-file load, only its part of the file is
    shipped, and the windows of its references refine to the whole-file results (region shards, no collective)."""
    from svtrek_b200 import api, shard as SH, simreads as S
    rs, truth = S.simulate(4, contig_lens=[250_000 if t % 5 else 200_000 for t in range(24)], coverage=4, read_len=8000, small_rate=0.01, n_sv=120, min_gap=1500)
    bam = str(tmp_path / "g.bam")
    S.write_bam(bam, rs, truth=truth)
    w = S.windows_from_truth(truth)
    a = api.Auditor(0)
    whole_st = a.load_bam(bam)
    whole = a.copy_reads()
    want = a.refine_windows(w["chrom"], w["start"], w["end"], w["pos"], w["kind"])
    shards = SH.plan_ref_shards(bam + ".bai", 3)
    assert len(shards) == 3
    got = np.full(len(want), -1, np.int32)
    for lo, hi in shards:
        st = a.load_bam(bam, refs=(lo, hi))
        part = a.copy_reads()
        r0, r1 = int(whole["tid_off"][lo]), int(whole["tid_off"][hi + 1])
        assert st["n_reads"] == r1 - r0 and st["image_bytes"] < 0.6 * whole_st["file_bytes"]
        for k in ("pos", "end", "n_cigar"):
            assert np.array_equal(part[k], whole[k][r0:r1]), k
        c0 = int(whole["cigar_off"][r0])
        assert np.array_equal(part["cigar"], whole["cigar"][c0:c0 + len(part["cigar"])])
        assert np.array_equal(part["tid_off"], np.clip(whole["tid_off"], r0, r1) - r0)
        mine = (w["chrom"] - 1 >= lo) & (w["chrom"] - 1 <= hi)
        res = a.refine_windows(w["chrom"], w["start"], w["end"], w["pos"], w["kind"])
        assert (res[~mine] == -1).all()
        got[mine] = res[mine]
    assert np.array_equal(got, want)
    # HTSlib-written file with an empty reference in the middle
    z = np.load(GOLD_NPZ)
    for lo, hi in ((0, 0), (1, 1), (2, 2), (0, 2), (1, 2)):
        st = a.load_bam(GOLD_BAM, refs=(lo, hi))
        part = a.copy_reads()
        r0, r1 = int(z["tid_off"][lo]), int(z["tid_off"][hi + 1])
        assert st["n_reads"] == r1 - r0
        assert np.array_equal(part["pos"], z["pos"][r0:r1]) and np.array_equal(part["end"], z["end"][r0:r1])
    with pytest.raises(api.SvtkError):
        a.load_bam(GOLD_BAM, refs=(2, 5))


@pytest.mark.gpu
def test_ranged_load_near_the_end_of_a_file_beyond_16_gib(tmp_path):
    """The decoder's 32-bit word index is relative to the bytes a load ships, so the 16 GiB limit applies to those, not to the file
    (ADVICE r1): a sparse 17 GiB BAM whose second reference sits at the very end loads by reference like a small file; only the
    whole-file load is refused.  (The hole is never read: a ranged load touches the header blocks and its own byte range.)"""
    import struct
    from oracle import binding as B
    from svtrek_b200 import api, simreads as S
    if S._packer() is None:
        pytest.skip("tools/libbgzfpack.so not built")
    L = 300_000
    parts, metas, sets = [], [], []
    for t in range(2):
        rs, truth = S.simulate(90 + t, contig_lens=[L], coverage=6, read_len=8000, small_rate=0.01, n_sv=30, min_gap=1500)
        p = str(tmp_path / f"part{t}.bam")
        m = S.write_bam_fast(p, rs, truth=truth, header_contigs=[L, L], tid_shift=t, want_meta=True)
        parts.append(p); metas.append(m); sets.append((rs, truth))
    hdr = int(metas[0]["hdr_bytes"])
    gap_at = 17 << 30                                          # where reference 1's blocks start
    big = str(tmp_path / "big.bam")
    try:
        with open(big, "wb") as f:
            with open(parts[0], "rb") as p0:
                f.write(p0.read(int(metas[0]["size"]) - 28))   # header blocks + reference 0's blocks
            f.seek(gap_at)                                     # a hole: no disk blocks behind it
            with open(parts[1], "rb") as p1:
                p1.seek(hdr)
                f.write(p1.read())                             # reference 1's blocks + the EOF block
        if os.stat(big).st_blocks * 512 > (1 << 30):
            pytest.skip("the file system does not keep the file sparse")
    except OSError as e:
        pytest.skip(f"cannot create a sparse 17 GiB file here: {e}")
    with open(big + ".bai", "wb") as f:
        f.write(b"BAI\1" + struct.pack("<i", 2))
        for t, shift in ((0, 0), (1, gap_at - hdr)):
            rs = sets[t][0]
            f.write(S.bai_section(metas[t]["bins"], metas[t]["v0"], metas[t]["v1"], rs.pos, rs.end, shift=shift))
    with api.Auditor(0) as a:
        with pytest.raises(api.SvtkError, match="16 GiB"):
            a.load_bam(big)
        for t in range(2):
            rs, truth = sets[t]
            st = a.load_bam(big, refs=(t, t))
            got = a.copy_reads()
            assert st["n_reads"] == rs.n_reads and st["image_bytes"] < 2 * os.path.getsize(parts[t])
            lo, hi = int(got["tid_off"][t]), int(got["tid_off"][t + 1])
            assert (lo, hi) == (0, rs.n_reads)
            for k, want in (("pos", rs.pos), ("end", rs.end), ("n_cigar", rs.n_cigar), ("cigar", rs.cigar)):
                assert np.array_equal(got[k].astype(np.int64), np.asarray(want).astype(np.int64)), (t, k)
            # the part's windows, asked for on reference t + 1 of the big file == the oracle on the part alone (its reference 1)
            w = S.windows_from_truth(truth)
            res = a.refine_windows(np.full_like(w["chrom"], t + 1), w["start"], w["end"], w["pos"], w["kind"])
            want = B.refine_batch(B.OracleReads(rs), w["chrom"], w["start"], w["end"], w["pos"], w["kind"])
            assert np.array_equal(res, want) and int((want >= 0).sum()) > 5, t
