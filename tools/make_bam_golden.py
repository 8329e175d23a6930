#!/usr/bin/env python
"""Writes tests/golden/bam_small.bam (+ .bai, + .npz): a small BAM produced by the REFERENCE's own HTSlib
(oracle/_ref/test_view, test_index -- built from /root/reference/htslib by oracle/build_ref.sh) from a SAM this script
generates, and the read arrays HTSlib renders from it.

The expected arrays come from tests/util_bam.py (plain Python: zlib + struct) and are cross-checked here against HTSlib's own
SAM rendering of the BAM (test_view BAM -> SAM): POS, CIGAR and FLAG of every record must agree, so the golden file is pinned
to HTSlib, not to the checker.  Cases inside: soft / hard clips, N, =, X, P ops; `*` CIGARs; FLAG 4 reads with a position;
a read with 70 000 CIGAR operations (stored by HTSlib as the `<l_seq>S<ref_len>N` placeholder + CG:B,I tag, sam.c:645); a fully
soft-clipped read that only looks like the placeholder; aux tags of every type; an empty reference in the middle; unplaced
reads at the end."""
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import binding as B          # noqa: E402
from tests import util_bam as U          # noqa: E402

OPS = "MIDNSHP=X"
GOLD = os.path.join(ROOT, "tests", "golden")


def rand_cigar(rng, target_q):
    """[(len, op)] with valid clipping structure, about target_q query bases"""
    core = []
    q = 0
    while q < target_q:
        op = rng.choice(["M", "M", "M", "=", "X", "I", "D", "N", "P"], p=[.4, .15, .15, .08, .04, .07, .07, .02, .02])
        ln = int(rng.integers(1, 120)) if op not in ("I", "D") or rng.random() < 0.8 else int(rng.integers(50, 400))
        if core and core[-1][1] == op:
            continue
        core.append((ln, op))
        if op in "MI=X":
            q += ln
    while core[0][1] not in "M=X":
        core.pop(0)
    while core[-1][1] not in "M=X":
        core.pop()
    pre, post = [], []
    if rng.random() < 0.3:
        pre.append((int(rng.integers(1, 300)), "S"))
    if rng.random() < 0.1:
        pre.insert(0, (int(rng.integers(1, 50)), "H"))
    if rng.random() < 0.3:
        post.append((int(rng.integers(1, 300)), "S"))
    if rng.random() < 0.1:
        post.append((int(rng.integers(1, 50)), "H"))
    return pre + core + post


def qlen(cig):
    return sum(l for l, o in cig if o in "MIS=X")


def main():
    rng = np.random.default_rng(20261104)
    refs = [("chr1", 200000), ("chr2", 50000), ("chr3", 300000)]
    recs = []   # (tid, pos0, flag, cigar list or None, aux string)
    auxes = ["", "NM:i:3", "NM:i:3\tMD:Z:10A5^AC6", "XA:A:Q\tXF:f:1.5\tXB:B:c,-1,2,3\tXS:B:S,1,65535\tXI:B:I,7,8\tXH:H:1AE301", "RG:Z:grp\tXC:i:-70000"]
    for tid, n in ((0, 160), (2, 90)):
        L = refs[tid][1]
        for p in np.sort(rng.integers(100, L - 20000, n)):
            cig = rand_cigar(rng, int(rng.integers(200, 3000)))
            flag = int(rng.choice([0, 16, 256, 2048, 1024, 0, 0]))
            recs.append((tid, int(p), flag, cig, auxes[int(rng.integers(0, len(auxes)))]))
    # special cases on chr1
    recs.append((0, 5000, 4, None, "XU:Z:placed_unmapped"))                       # FLAG 4 with a position, CIGAR *
    recs.append((0, 5000, 0, None, ""))                                             # mapped, CIGAR *
    recs.append((0, 6000, 4, [(150, "M")], ""))                                     # FLAG 4 but a CIGAR: endpos is pos + 1
    recs.append((0, 7000, 0, [(100, "S")], "XZ:Z:all_clipped\tXB:B:i,1,2"))          # looks like the CG placeholder, no CG tag
    recs.append((0, 7000, 0, [(100, "S")], "CG:Z:not_an_array"))                     # CG tag of the wrong type
    long_cig = []
    for k in range(35000):
        long_cig += [(1, "M"), (1, "I")]
    long_cig.append((3, "M"))
    recs.append((0, 90000, 0, long_cig, "NM:i:35000"))                               # 70 001 ops: CG:B,I on disk
    recs.sort(key=lambda r: (r[0], r[1]))
    sam = os.path.join(GOLD, "_bam_small.sam")
    with open(sam, "w") as f:
        f.write("@HD\tVN:1.6\tSO:coordinate\n")
        for n, L in refs:
            f.write(f"@SQ\tSN:{n}\tLN:{L}\n")
        f.write("@CO\t" + "padding " * 3000 + "\n")          # a header longer than one would expect in tests
        for i, (tid, p, flag, cig, aux) in enumerate(recs):
            q = qlen(cig) if cig else int(rng.integers(20, 200))
            seq = "".join(rng.choice(list("ACGT"), q)) if q else "*"
            qual = "".join(chr(33 + int(x)) for x in rng.integers(10, 41, q)) if q else "*"
            cs = "".join(f"{l}{o}" for l, o in cig) if cig else "*"
            f.write(f"r{i}\t{flag}\t{refs[tid][0]}\t{p + 1}\t60\t{cs}\t*\t0\t0\t{seq}\t{qual}" + ("\t" + aux if aux else "") + "\n")
        for i in range(12):                                                            # unplaced reads: refID -1, last in the file
            q = int(rng.integers(50, 3000))
            f.write(f"u{i}\t4\t*\t0\t0\t*\t*\t0\t0\t{''.join(rng.choice(list('ACGT'), q))}\t{'I' * q}\n")
    bam = os.path.join(GOLD, "bam_small.bam")
    B.make_bam(sam, bam)
    os.remove(sam)
    got = U.read_bam(bam)
    # pin the checker to HTSlib: its SAM rendering of the same BAM
    txt = subprocess.run([os.path.join(B.REF_DIR, "test_view"), bam], stdout=subprocess.PIPE, check=True).stdout.decode().splitlines()
    rows = [l.split("\t") for l in txt if not l.startswith("@")]
    placed = [r for r in rows if r[2] != "*"]
    assert len(rows) == got["n_records"] and len(placed) == len(got["pos"]), (len(rows), got["n_records"], len(placed))
    for k, r in enumerate(placed):
        w = got["cigar"][int(got["cigar_off"][k]):int(got["cigar_off"][k]) + int(got["n_cigar"][k])]
        cs = "".join(f"{int(x) >> 4}{OPS[int(x) & 15]}" for x in w) or "*"
        assert int(r[3]) - 1 == int(got["pos"][k]) and r[5] == cs, (k, r[3], r[5][:40], cs[:40])
    assert int(got["n_cigar"].max()) == 70001
    np.savez_compressed(os.path.join(GOLD, "bam_small.npz"), **{k: got[k] for k in ("tid_off", "pos", "end", "n_cigar", "cigar_off", "cigar")},
                        ref_names=np.array([n for n, _ in got["refs"]]), ref_lens=np.array([l for _, l in got["refs"]], np.int64))
    print("wrote", bam, os.path.getsize(bam), "bytes;", len(got["pos"]), "placed reads of", got["n_records"], "records;",
          len(U.bgzf_blocks(open(bam, "rb").read())), "BGZF blocks")


if __name__ == "__main__":
    main()
