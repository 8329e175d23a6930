#!/usr/bin/env python
"""Regenerate tests/golden/ by RUNNING THE UNMODIFIED REFERENCE built under oracle/_ref (oracle/build_ref.sh).

Run in the build container only (needs /root/reference to have been compiled):  python tools/make_golden.py
The GPU box has no /root/reference, so the vectors this writes are committed; tests read them from disk.
Every expected value in tests/golden/ comes from the reference's own code (libsvtrek_ref.so direct calls
or the `svtrek audt -t 1` binary on a padded VCF) -- none from the oracle or the CUDA path."""
import json
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import binding as B  # noqa: E402
from svtrek_b200 import simreads as S  # noqa: E402

GOLD = os.environ.get("SVTK_GOLDEN_OUT") or os.path.join(ROOT, "tests", "golden")      # SVTK_GOLDEN_OUT: tests/test_golden_drift.py regenerates into a scratch directory
os.makedirs(GOLD, exist_ok=True)


def save_reads(path, rs, **extra):
    np.savez_compressed(path, tid_off=rs.tid_off, pos=rs.pos, end=rs.end, n_cigar=rs.n_cigar, cigar_off=rs.cigar_off,
                        cigar=rs.cigar, contig_len=np.array(rs.contig_len, dtype=np.int64), **extra)


# ------------------------------------------------------------------------------------------------ A
def gen_consensus_cases(rng, n_cases):
    cases = []
    fixed = [([100, 101, 102, 480, 481, 482, 483, 900], 490, 3, 5, 500), ([100, 101, 102, 480, 481, 482, 483, 900], 110, 3, 5, 500),
             ([600, 601, 602], 560, 3, 5, 500), ([10, 600, 601, 602], 560, 3, 5, 500), ([], 100, 3, 5, 500), ([5], 5, 1, 5, 500),
             ([5, 5, 5], 5, 3, 0, 500), ([1, 2, 3], 1000, 1, 5, 500), ([100] * 10 + [130] * 12, 115, 3, 5, 500),
             ([100] * 10 + [130] * 12, 115, 3, 5, 10), ([90, 95, 100, 105, 110, 115], 100, 2, 5, 500),
             ([2147483000, 2147483001, 2147483002], 2147483001, 3, 5, 500)]
    cases += fixed
    for _ in range(n_cases):
        pos = int(rng.integers(0, 3_000_000))
        style = rng.integers(0, 6)
        n = int(rng.integers(0, 40)) if style < 4 else int(rng.integers(40, 700))
        vals = []
        n_cl = int(rng.integers(1, 5))
        centers = pos + rng.integers(-700, 700, n_cl)
        spread = int(rng.choice([0, 1, 2, 3, 6, 12, 40]))
        for i in range(n):
            if rng.random() < 0.15:
                vals.append(int(pos + rng.integers(-20000, 20000)))
            else:
                c = int(centers[rng.integers(0, n_cl)])
                vals.append(c + int(rng.integers(-spread, spread + 1)))
        vals = [max(0, v) for v in vals]
        minc = int(rng.choice([1, 2, 3, 3, 3, 5, 10]))
        interval = int(rng.choice([0, 1, 5, 5, 5, 20, 60]))
        rng_ = int(rng.choice([10, 100, 500, 500, 500, 2000]))
        cases.append((vals, pos, minc, interval, rng_))
    return cases


def golden_consensus():
    rng = np.random.default_rng(11)
    cases = gen_consensus_cases(rng, 3000)
    vals, off, meta, exp, lb, ub = [], [0], [], [], [], []
    for v, pos, minc, interval, rg in cases:
        r = B.ref_consensus_pos(v, pos, minc, interval, rg)
        o = B.consensus_pos(v, pos, minc, interval, rg)
        assert r == o, ("oracle != reference", v, pos, minc, interval, rg, r, o)
        vals += v; off.append(len(vals)); meta.append((pos, minc, interval, rg)); exp.append(r)
        sv = np.sort(np.array(v, dtype=np.int32))
        if len(v):
            a = np.ascontiguousarray(sv)
            l_ref = B.ref().ref_lower_bound(a.ctypes.data, len(v), pos + 25)
            u_ref = B.ref().ref_upper_bound(a.ctypes.data, len(v), pos - 25)
            assert l_ref == B.lower_bound(a, pos + 25) and u_ref == B.upper_bound(a, pos - 25)
        else:
            l_ref = u_ref = -1
        lb.append(l_ref); ub.append(u_ref)
    np.savez_compressed(os.path.join(GOLD, "consensus_pos.npz"), values=np.array(vals, dtype=np.int32),
                        offsets=np.array(off, dtype=np.int64), meta=np.array(meta, dtype=np.int32),
                        expected=np.array(exp, dtype=np.int32), lower=np.array(lb, dtype=np.int32), upper=np.array(ub, dtype=np.int32))
    print(f"consensus_pos: {len(cases)} cases, {sum(1 for e in exp if e != -1)} non-NA; oracle == reference")


# ------------------------------------------------------------------------------------------------ B / C
def parse_cigar(s):
    out, num = [], 0
    for ch in s:
        if ch.isdigit():
            num = num * 10 + int(ch)
        else:
            out.append(num << 4 | S.OPCHARS.index(ch)); num = 0
    return out


def readset_from_list(reads, contig_lens, flags=None):
    """reads: list of (tid, pos0, cigar_string) -> ReadSet sorted by (tid,pos) (stable)."""
    idx = sorted(range(len(reads)), key=lambda i: (reads[i][0], reads[i][1]))
    reads = [reads[i] for i in idx]
    flags = None if flags is None else [flags[i] for i in idx]
    words, off, pos, ncig, tid_off = [], [0], [], [], [0] * (len(contig_lens) + 1)
    for t, p, c in reads:
        w = parse_cigar(c)
        words += w; off.append(len(words)); pos.append(p); ncig.append(len(w)); tid_off[t + 1] += 1
    tid_off = np.cumsum(tid_off).astype(np.int64)
    rs = S.ReadSet(tid_off, np.array(pos, dtype=np.uint32), None, np.array(ncig, dtype=np.uint32),
                   np.array(off, dtype=np.uint64), np.array(words, dtype=np.uint32), list(contig_lens))
    rs.end = S.endpos_from_cigar(rs.pos, rs.cigar_off, rs.cigar)
    if flags is not None:
        fl = np.array(flags, dtype=np.int64)
        rs.end = np.where((fl & 4) != 0, rs.pos + 1, rs.end).astype(np.uint32)   # bam_endpos: BAM_FUNMAP => length 1
    return rs, flags


def to_bam(rs, tmp, name, flags=None, random_seq=False):
    sam = os.path.join(tmp, name + ".sam"); bam = os.path.join(tmp, name + ".bam")
    S.write_sam(sam, rs, flags=flags, random_seq=random_seq)
    B.make_bam(sam, bam)
    return bam


QUIRK_READS = [  # (tid, pos0, cigar) -- one read per contig: SURVEY.md G12-G22
    (0, 1000, "5H100M60D100M"), (1, 1000, "100M5P60D100M"), (2, 1000, "100M60N60D100M"), (3, 1000, "100M100D100M"),
    (4, 1000, "100M60D5000M"), (5, 2000, "50S300M"), (6, 1000, "500M50S"), (7, 1000, "100M50I100M"), (8, 1000, "100M49I100M"),
    (9, 1000, "100M50D100M"), (10, 1000, "100M51D100M"), (11, 1000, "20S100M70D100M30S"), (12, 1000, "100M70D100=5X60I10M")]
QUIRK_QUERIES = [  # (kind, chrom, start, end)
    (0, 1, 1000, 1300), (0, 2, 1000, 1300), (0, 3, 1000, 1400), (0, 4, 1000, 1150), (1, 4, 1000, 1150), (0, 5, 5000, 6000),
    (1, 6, 1900, 2100), (0, 7, 1400, 1600), (0, 7, 1400, 1450), (2, 8, 900, 1300), (2, 9, 900, 1300), (0, 10, 900, 1300),
    (1, 10, 900, 1300), (0, 11, 900, 1300), (1, 11, 900, 1300), (0, 1, 0, 1300), (0, 99, 900, 1300), (0, 0, 900, 1300),
    (3, 11, 900, 1300), (0, 12, 900, 1400), (1, 12, 900, 1400), (2, 12, 900, 1400), (0, 13, 900, 1500), (1, 13, 900, 1500),
    (2, 13, 900, 1500), (2, 13, 900, 1250), (0, 12, 1250, 1400), (1, 12, 1100, 1200), (0, 1, 1300, 1000), (2, 8, 1101, 1102),
    (2, 8, 1, 1000), (2, 8, 1, 1001), (2, 8, 1201, 1300), (2, 8, 1200, 1300), (0, 4, 4294967000, 1150)]


def golden_quirks(tmp):
    rs, _ = readset_from_list(QUIRK_READS, [20000] * 13)
    bam = to_bam(rs, tmp, "quirk")
    rb = B.RefBam(bam); orc = B.OracleReads(rs)
    out = []
    for kind, chrom, st, en in QUIRK_QUERIES:
        r, sig = rb.refine(kind, chrom, st, en, 1100, 1, 5, 1_000_000, want_sig=True)
        o, osig = B.refine(orc, kind, chrom, st, en, 1100, 1, 5, 1_000_000, want_sig=True)
        assert r == o and np.array_equal(sig, osig), ("quirk mismatch", kind, chrom, st, en, r, o, sig, osig)
        out.append(dict(kind=kind, chrom=chrom, start=st, end=en, pos=1100, min_count=1, interval=5, range=1_000_000,
                        expected=int(r), signatures=[int(x) for x in sig]))
    rb.close()
    json.dump(dict(reads=QUIRK_READS, contig_len=[20000] * 13, queries=out), open(os.path.join(GOLD, "walk_quirks.json"), "w"), indent=0)
    print(f"walk_quirks: {len(out)} queries; oracle == reference")


def gen_mix_reads(rng, n_reads, contig_lens):
    reads, flags = [], []
    hot = [np.sort(rng.integers(2000, L - 2000, 25)) for L in contig_lens]
    for _ in range(n_reads):
        tid = int(rng.choice([0, 0, 0, 1, 1, 3]))      # contig 2 (index) stays empty
        L = contig_lens[tid]
        pos = int(rng.integers(0, L - 100))
        nops = int(rng.integers(1, 28))
        cig, ref = [], pos
        if rng.random() < 0.1: cig.append(f"{int(rng.integers(1, 50))}H")
        if rng.random() < 0.3: cig.append(f"{int(rng.integers(1, 600))}S")
        for _k in range(nops):
            m = int(rng.integers(1, 900))
            h = hot[tid]
            j = np.searchsorted(h, ref)
            if j < len(h) and h[j] < ref + m and rng.random() < 0.7 and h[j] > ref:   # land exactly on a hot breakpoint
                m = int(h[j] - ref)
                ev = rng.choice(["D", "I"])
                ln = int(rng.choice([50, 51, 60, 120, 300]))
                cig.append(f"{m}{rng.choice(['M', 'M', '=', 'X'])}"); ref += m
                cig.append(f"{ln}{ev}"); ref += ln if ev == "D" else 0
                continue
            cig.append(f"{m}{rng.choice(['M', 'M', 'M', '=', 'X'])}"); ref += m
            u = rng.random()
            if u < 0.25: cig.append(f"{int(rng.integers(1, 6))}I")
            elif u < 0.5: ln = int(rng.integers(1, 6)); cig.append(f"{ln}D"); ref += ln
            elif u < 0.58: ln = int(rng.choice([49, 50, 51, 52, 200])); cig.append(f"{ln}D"); ref += ln
            elif u < 0.66: cig.append(f"{int(rng.choice([49, 50, 51, 400]))}I")
            elif u < 0.70: ln = int(rng.integers(10, 700)); cig.append(f"{ln}N"); ref += ln
            elif u < 0.72: cig.append(f"{int(rng.integers(1, 9))}P")
            elif u < 0.735: cig.append(f"{int(rng.integers(1, 90))}S")     # interior clip: exercises the op rule only
            elif u < 0.745: cig.append(f"{int(rng.integers(1, 90))}H")
        if rng.random() < 0.6: cig.append(f"{int(rng.integers(1, 300))}M")
        if rng.random() < 0.3: cig.append(f"{int(rng.integers(1, 600))}S")
        if rng.random() < 0.1: cig.append(f"{int(rng.integers(1, 50))}H")
        reads.append((tid, pos, "".join(cig)))
        flags.append(int(rng.choice([0, 0, 0, 16, 256, 1024, 2048, 4])))
    return reads, flags, hot


def golden_mix(tmp):
    rng = np.random.default_rng(5)
    contig_lens = [120_000, 60_000, 50_000, 30_000]
    reads, flags, hot = gen_mix_reads(rng, 1500, contig_lens)
    rs, flags = readset_from_list(reads, contig_lens, flags)
    bam = to_bam(rs, tmp, "mix", flags=flags)
    rb = B.RefBam(bam); orc = B.OracleReads(rs)
    q, exp, sigs, soff = [], [], [], [0]
    n_non_na = 0
    for i in range(4000):
        kind = int(rng.integers(0, 4))
        chrom = int(rng.choice([1, 1, 1, 2, 2, 4, 3, 0, 5, 9]))
        L = contig_lens[chrom - 1] if 1 <= chrom <= 4 else 50_000
        h = hot[chrom - 1] if 1 <= chrom <= 4 and chrom != 3 else np.array([1000])
        pos = int(h[rng.integers(0, len(h))] + rng.integers(-60, 60)) if rng.random() < 0.8 else int(rng.integers(0, L))
        half = int(rng.choice([50, 300, 2000, 10000, 20000]))
        st, en = (pos - half) & 0xFFFFFFFF, (pos + int(rng.choice([half, 2000]))) & 0xFFFFFFFF
        if rng.random() < 0.03: st = 0
        if rng.random() < 0.02: en = 0
        minc = int(rng.choice([1, 2, 3, 3])); interval = int(rng.choice([1, 5, 5, 20])); rg = int(rng.choice([100, 500, 500, 3000]))
        r, sig = rb.refine(kind, chrom, st, en, pos, minc, interval, rg, want_sig=True)
        o, osig = B.refine(orc, kind, chrom, st, en, pos, minc, interval, rg, want_sig=True)
        assert r == o and np.array_equal(sig, osig), ("mix mismatch", i, kind, chrom, st, en, pos, r, o, len(sig), len(osig))
        n_non_na += r != -1
        q.append((kind, chrom, st, en, pos, minc, interval, rg)); exp.append(r); sigs += [int(x) for x in sig]; soff.append(len(sigs))
    # sliding_window_ins direct calls (dead code in the reference, reachable only this way)
    sw, sw_lines = [], []
    for i in range(60):
        chrom = int(rng.choice([1, 1, 2, 4]))
        L = contig_lens[chrom - 1]
        st = int(rng.integers(1, L - 5000)); en = st + int(rng.choice([1000, 5000, 20000]))
        win = int(rng.choice([100, 500, 2000])); slide = int(rng.choice([1, 1, 3])); minc = int(rng.choice([1, 2, 3]))
        r, rl = rb.sliding_window_ins(chrom, st, en, minc, win, slide, want_lines=True)
        o, ol = B.sliding_window_ins(orc, chrom, st, en, minc, win, slide, want_lines=True)
        assert r == o and rl == ol, ("sliding mismatch", chrom, st, en, minc, win, slide, r, o)
        sw.append((chrom, st, en, minc, win, slide, r)); sw_lines.append(rl)
    rb.close()
    save_reads(os.path.join(GOLD, "mix_reads.npz"), rs, flags=np.array(flags, dtype=np.int32),
               queries=np.array(q, dtype=np.int64), expected=np.array(exp, dtype=np.int32),
               sig=np.array(sigs, dtype=np.int32), sig_off=np.array(soff, dtype=np.int64), sliding=np.array(sw, dtype=np.int64))
    json.dump(sw_lines, open(os.path.join(GOLD, "mix_sliding_lines.json"), "w"), indent=0)
    print(f"mix: {len(q)} windows ({n_non_na} non-NA, {soff[-1]} signatures), {len(sw)} sliding-window calls; oracle == reference")


# ------------------------------------------------------------------------------------------------ D
EDGE_LINES = [  # parser edge cases (SURVEY.md section 8c "VCF-parser vectors"); positions refer to the c1mini set
    "chrX\t{p}\t.\tN\tN\t.\tPASS\tSVTYPE=DEL;END={e}",
    "chr1\t{p}\t.\tN\t" + "A" * 60 + "\t.\tPASS\tFOO=1",
    "chr1\t{p}\t.\t" + "A" * 101 + "\tA\t.\tPASS\tFOO=1",
    "chr1\t{p}\t.\tN\tN\t.\tPASS\tSVTYPE=DEL;END={p50}",
    "chr1\t{p}\t.\tN\tN\t.\tPASS\tSVTYPE=DEL;END={p49}",
    "chr1\t{p}\t.\tN\tN\t.\tPASS\tSVTYPE=DEL;END={p51}",
    "chr1\t{p}\t.\tN\tN\t.\tPASS\tSVTYPE=INV;END={e}",
    "chr1\t{p}\t.\tN\tN\t.\tPASS\tSVTYPE=DUP;END={e}",
    "chr1\t{p}\t.\tN\tN\t.\tPASS\tSVTYPE=DELETION;END={e}",
    "chr1\t{p}\t.\tN\t<INS:ME>\t.\tPASS\tSVTYPE=INS:ME",
    "chr1\t{p}\t.\tN\t<DEL:ME>\t.\tPASS\tEND={e};SVTYPE=DEL:ME",
    "chr1\t{e}\t.\tN\tN\t.\tPASS\tSVTYPE=DEL;END={p}",
    "chr1_KI270706v1_random\t{p}\t.\tN\tN\t.\tPASS\tEND={e};SVTYPE=DEL",
    "chr1\t{p}\t.\tA\tC,G\t.\tPASS\tFOO=1",
    "1\t{p}\t.\tN\tN\t.\tPASS\tEND={e};SVTYPE=DEL",
    "chr1\t5000\t.\tN\tN\t.\tPASS\tEND=5400;SVTYPE=DEL",
    "chr1\tabc\t.\tN\tN\t.\tPASS\tEND=5400;SVTYPE=DEL",
    "chr1\t0\t.\tN\tN\t.\tPASS\tSVTYPE=INS",
    "chr1\t{p}\t.\tN\tN\t.\tPASS\tEND=xyz;SVTYPE=DEL",
    "chr1\t{p}\t.\tN\tN\t.\tPASS\tSVTYPE=BND",
    "chr2\t{p}\t.\tN\tN\t.\tPASS\tEND={e};SVTYPE=DEL",
    "chr1\t{p}\t.\tN\tN\t.\tPASS\tSVTYPE=INS;SOMEEND=12;X",
    "lonely",
    "chr1\t{p}\t.\tN\tN\t.\tPASS\tCIEND=-10,10;END={e};SVTYPE=DEL",
    "chr1\t\t{p}\t.\tN\tN\t.\tPASS\tEND={e};SVTYPE=DEL",
    "chr1\t{p}\t.\tN\tN\t.\tPASS\tSVTYPE=INS;END={p}\tGT\t0/1",
]


def golden_cli(tmp):
    rs, truth = S.simulate(424242, [260_000], 30, 15000, 0.005, 90, min_gap=1500, margin=30000)
    bam = to_bam(rs, tmp, "c1mini", random_seq=False)
    orc = B.OracleReads(rs)
    d0 = int(truth.pos0[np.flatnonzero(truth.is_del)[3]]); l0 = int(truth.length[np.flatnonzero(truth.is_del)[3]])
    fmt = dict(p=d0 + 7, e=d0 + l0 - 4, p50=d0 + 7 + 50, p49=d0 + 7 + 49, p51=d0 + 7 + 51)
    variants = {
        "simvcf": S.vcf_lines(truth, seed=3, end_first=False),
        "endfirst": S.vcf_lines(truth, seed=3, end_first=True),
        "edge": [ln.format(**fmt) for ln in EDGE_LINES],
    }
    sweeps = {"default": [], "i1_c2": ["--consensus-interval", 1, "--consensus-min-count", 2],
              "i20_c10": ["--consensus-interval", 20, "--consensus-min-count", 10],
              "narrowwin": ["--wider-interval", 3000, "--median-interval", 1500, "--narrow-interval", 400, "--consensus-interval-range", 60]}
    key2param = {"--consensus-interval": "interval", "--consensus-min-count": "min_count", "--wider-interval": "wider",
                 "--median-interval": "median", "--narrow-interval": "narrow", "--consensus-interval-range": "range"}
    gold = {}
    for vname, lines in variants.items():
        vcf = os.path.join(tmp, vname + ".vcf")
        S.write_vcf(vcf, lines, pad=16)           # reference drops up to 2t queued lines at EOF (audit.c:35-41,342)
        plain = os.path.join(tmp, vname + ".plain.vcf")
        S.write_vcf(plain, lines, pad=0)
        for sname, extra in sweeps.items():
            if vname == "edge" and sname != "default":
                continue
            out, err = B.run_ref_cli(bam, vcf, threads=1, extra=extra)
            params = {key2param[extra[i]]: extra[i + 1] for i in range(0, len(extra), 2)}
            o_out, o_err = B.audit_file(orc, plain, params, want_err=True)
            err_nopad = "".join(l + "\n" for l in err.splitlines() if l != "VCF: no index at line: pad")
            assert out == o_out, ("CLI stdout mismatch", vname, sname, _first_diff(out, o_out))
            assert err_nopad == o_err, ("CLI stderr mismatch", vname, sname, err_nopad, o_err)
            gold[f"{vname}/{sname}"] = dict(params=params, stdout=out, stderr=err_nopad)
        gold.setdefault("_vcf", {})[vname] = lines
    save_reads(os.path.join(GOLD, "c1mini_reads.npz"), rs)
    json.dump(gold, open(os.path.join(GOLD, "c1mini_cli.json"), "w"), indent=0)
    n_lines = sum(v["stdout"].count("\n") - 2 for k, v in gold.items() if k != "_vcf")
    print(f"cli: {len(gold) - 1} runs of `svtrek audt -t 1`, {n_lines} record lines; oracle stdout+stderr == reference")


def _first_diff(a, b):
    for i, (x, y) in enumerate(zip(a.splitlines(), b.splitlines())):
        if x != y:
            return i, x, y
    return "length", len(a.splitlines()), len(b.splitlines())


if __name__ == "__main__":
    assert B.have_ref(), "oracle/_ref missing: run `make -C oracle ref` in the build container first"
    with tempfile.TemporaryDirectory() as tmp:
        golden_consensus()
        golden_quirks(tmp)
        golden_mix(tmp)
        if os.environ.get("SVTK_GOLDEN_SKIP_CLI") != "1":      # the CLI captures belong to the stored c1mini read set (c1mini_reads.npz), not to a fresh simulate()
            golden_cli(tmp)
    print("golden vectors written to", GOLD)
