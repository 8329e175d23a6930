#!/usr/bin/env python
"""tests/golden/ins_sequences.npz: the inserted sequences behind refined insertion breakpoints, extracted with the REFERENCE'S
HTSlib (oracle/ins_harness.c: sam_itr_queryi + sam_itr_next + refine_ins' loop + bam_seqi) from a small seeded BAM.  The BAM itself
is not stored: tests rebuild it with ins_case() below (same seeds -> same records) and compare svtk_ins_sequences with the file.
Run here (needs oracle/_ref); the GPU box only reads the .npz."""
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
OUT = os.path.join(ROOT, "tests", "golden", "ins_sequences.npz")


def ins_case(path):
    """The fixture's inputs: a 250 kb contig, 25x 9 kb reads with 1 % small indels, 40 INS / DEL, carriers of an insertion hold a
    mutated copy of it; BAM written to `path`.  Returns (read set, truth, windows, refined-by-the-oracle)."""
    from oracle import binding as B
    from svtrek_b200 import simreads as S
    rs, truth = S.simulate(4242, [250_000], 25, 9000, 0.01, 40, margin=12000, min_gap=900, max_len=700)
    S.write_bam(path, rs, seed=11, truth=truth)
    w = S.windows_from_truth(truth, seed=3)
    refined = B.refine_batch(B.OracleReads(rs), w["chrom"], w["start"], w["end"], w["pos"], w["kind"])
    return rs, truth, w, refined


def main():
    from oracle import binding as B
    assert B.have_ins_ref(), "run oracle/build_ref.sh first"
    with tempfile.TemporaryDirectory() as d:
        bam = os.path.join(d, "ins.bam")
        rs, truth, w, refined = ins_case(bam)
        cases = {}
        for tol, min_len in ((5, 50), (0, 50), (40, 120)):
            r = B.ref_ins_sequences(bam, w["chrom"], w["start"], w["end"], w["kind"], refined, tol=tol, min_len=min_len)
            for k, v in r.items():
                cases[f"t{tol}_m{min_len}_{k}"] = v
            print(f"tol {tol} min_len {min_len}: {len(r['window'])} sequences in {len(np.unique(r['window']))} windows, {len(r['bases'])} bases")
    np.savez_compressed(OUT, refined=refined, n_reads=rs.n_reads, n_win=len(w["chrom"]), **cases)
    print("wrote", OUT, os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
    main()
