#!/usr/bin/env python
"""Writes tests/golden/poa_cases.npz from the compiled, unmodified abPOA library (oracle/_ref/libabpoa_ref.so, built
by oracle/build_ref.sh from /root/reference/abPOA).  Run in the build container (needs /root/reference for the three
abPOA test_data FASTA files); the .npz travels, /root/reference does not.

Per read set: the input reads and abpoa's consensus, coverage, and per read best_score, graph size after the read,
number of alignment ops and their FNV-1a hash."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import binding as B          # noqa: E402
from svtrek_b200 import simpoa as P      # noqa: E402


def read_fa(p):
    seqs, cur = [], []
    for ln in open(p):
        ln = ln.strip()
        if ln.startswith(">"):
            if cur:
                seqs.append("".join(cur)); cur = []
        elif ln:
            cur.append(ln)
    if cur:
        seqs.append("".join(cur))
    return [P.encode(s) for s in seqs]


def main():
    sets, names = [], []
    td = "/root/reference/abPOA/test_data"
    for f in ("seq.fa", "test.fa", "heter.fa"):
        sets.append(read_fa(os.path.join(td, f))); names.append("abpoa:" + f)
    rng = np.random.default_rng(20261104)
    for t in range(45):
        kind = ("random", "tandem", "ragged")[t % 3]
        s = P.read_set(rng, kind, length=int(rng.integers(50, 700)) if kind != "tandem" else None)
        if kind == "tandem":
            s = s[:24]
        sets.append(s); names.append(f"{kind}:{t}")
    # edge cases: one read, two identical reads, unrelated reads (graph ~ sum of lengths), all-N read, length-1 reads
    sets.append([rng.integers(0, 4, 80).astype(np.uint8)]); names.append("edge:single")
    x = rng.integers(0, 4, 120).astype(np.uint8)
    sets.append([x, x.copy()]); names.append("edge:identical")
    sets.append([rng.integers(0, 4, 150).astype(np.uint8) for _ in range(12)]); names.append("edge:unrelated")
    sets.append([x, np.full(100, 4, np.uint8), x.copy()]); names.append("edge:allN")
    sets.append([np.array([2], np.uint8), np.array([2], np.uint8), np.array([1], np.uint8)]); names.append("edge:len1")
    sets.append([x, x[:10].copy(), x[60:].copy(), x.copy()]); names.append("edge:fragments")
    set_off, seq_off, bases = P.pack(sets)
    cons, cov, bs, nn, no, hh, cc = [], [], [], [], [], [], []
    for s in sets:
        r = B.ref_poa_steps(s)
        m, _ = B.ref_poa_msa(s)
        assert np.array_equal(m, r["cons"]), "abpoa_msa and the unrolled path disagree"
        cons.append(r["cons"]); cov.append(r["cov"]); bs.append(r["best_score"]); nn.append(r["node_n"]); no.append(r["n_ops"])
        hh.append(r["ops_hash"]); cc.append(r["cells"])
    cons_off = np.zeros(len(sets) + 1, np.int64); cons_off[1:] = np.cumsum([len(c) for c in cons])
    out = os.path.join(ROOT, "tests", "golden", "poa_cases.npz")
    np.savez_compressed(out, names=np.array(names), set_off=set_off, seq_off=seq_off, bases=bases, cons_off=cons_off,
                        cons=np.concatenate(cons), cov=np.concatenate(cov), best_score=np.concatenate(bs), node_n=np.concatenate(nn),
                        n_ops=np.concatenate(no), ops_hash=np.concatenate(hh), cells=np.concatenate(cc))
    print(out, os.path.getsize(out), "bytes;", len(sets), "read sets,", int(seq_off[-1]), "bases")
    for i in (0, 1):
        print(names[i], P.decode(cons[i]))


if __name__ == "__main__":
    main()
