"""The CPU oracle (oracle/svtk_oracle.c) against vectors captured from the UNMODIFIED reference
(tests/golden/, written by tools/make_golden.py from oracle/_ref).  Runs without a GPU."""
import os
import tempfile

import numpy as np

from oracle import binding as B
from svtrek_b200 import simreads as S
from tests import util


def test_consensus_pos_golden():
    z = np.load(os.path.join(util.GOLD, "consensus_pos.npz"))
    vals, off, meta, exp = z["values"], z["offsets"], z["meta"], z["expected"]
    assert len(exp) > 3000
    for i in range(len(exp)):
        v = vals[off[i]:off[i + 1]]
        pos, minc, interval, rg = (int(x) for x in meta[i])
        assert B.consensus_pos(v, pos, minc, interval, rg) == exp[i], i
        if len(v):
            sv = np.sort(v)
            assert B.lower_bound(sv, pos + 25) == z["lower"][i]
            assert B.upper_bound(sv, pos - 25) == z["upper"][i]


def test_walk_quirks_golden():
    rs, queries = util.quirk_readset()
    orc = B.OracleReads(rs)
    for q in queries:
        r, sig = B.refine(orc, q["kind"], q["chrom"], q["start"], q["end"], q["pos"], q["min_count"], q["interval"], q["range"], want_sig=True)
        assert r == q["expected"], q
        assert [int(x) for x in sig] == q["signatures"], q


def test_mix_windows_golden():
    rs, z = util.load_reads("mix_reads.npz")
    orc = B.OracleReads(rs)
    q, exp, sig, soff = z["queries"], z["expected"], z["sig"], z["sig_off"]
    for i in range(len(exp)):
        kind, chrom, st, en, pos, minc, interval, rg = (int(x) for x in q[i])
        r, s = B.refine(orc, kind, chrom, st, en, pos, minc, interval, rg, want_sig=True)
        assert r == exp[i], i
        assert np.array_equal(s, sig[soff[i]:soff[i + 1]]), i


def test_sliding_window_golden():
    rs, z = util.load_reads("mix_reads.npz")
    orc = B.OracleReads(rs)
    lines = util.load_json("mix_sliding_lines.json")
    for i, row in enumerate(z["sliding"]):
        chrom, st, en, minc, win, slide, exp = (int(x) for x in row)
        r, ln = B.sliding_window_ins(orc, chrom, st, en, minc, win, slide, want_lines=True)
        assert r == exp and ln == lines[i], i


def test_cli_stdout_golden():
    rs, _ = util.load_reads("c1mini_reads.npz")
    orc = B.OracleReads(rs)
    gold = util.load_json("c1mini_cli.json")
    keymap = dict(interval="interval", min_count="min_count", wider="wider", median="median", narrow="narrow", range="range")
    with tempfile.TemporaryDirectory() as d:
        for key, g in gold.items():
            if key == "_vcf":
                continue
            vname = key.split("/")[0]
            vcf = os.path.join(d, vname + ".vcf")
            S.write_vcf(vcf, gold["_vcf"][vname])
            out, err = B.audit_file(orc, vcf, {keymap[k]: v for k, v in g["params"].items()}, want_err=True)
            assert out == g["stdout"], key
            assert err == g["stderr"], key
