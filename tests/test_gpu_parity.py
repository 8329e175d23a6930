"""GPU parity tests proper: the CUDA path, called through the C ABI (libsvtk_b200.so via svtrek_b200.api),
against (a) vectors captured from the unmodified reference (tests/golden) and (b) the pinned CPU oracle on
seeded synthetic inputs.  Bit-exact: everything on this path is integer / index work."""
import os
import tempfile

import numpy as np
import pytest

from oracle import binding as B
from svtrek_b200 import api
from svtrek_b200 import simreads as S
from tests import util

pytestmark = pytest.mark.gpu


def test_consensus_pos_golden(auditor):
    z = np.load(os.path.join(util.GOLD, "consensus_pos.npz"))
    vals, off, meta, exp = z["values"], z["offsets"], z["meta"], z["expected"]
    groups = {}
    for i in range(len(exp)):
        groups.setdefault(tuple(int(x) for x in meta[i][1:]), []).append(i)
    n_checked = 0
    for (minc, interval, rg), idx in groups.items():
        auditor.set_params(consensus_min_count=minc, consensus_interval=interval, consensus_interval_range=rg)
        v = np.concatenate([vals[off[i]:off[i + 1]] for i in idx]) if idx else np.zeros(0, np.int32)
        o = np.concatenate(([0], np.cumsum([off[i + 1] - off[i] for i in idx])))
        got = auditor.consensus_pos_batch(v, o, meta[idx, 0])
        assert np.array_equal(got, exp[idx]), (minc, interval, rg, np.flatnonzero(got != exp[idx])[:5])
        n_checked += len(idx)
    assert n_checked == len(exp)
    auditor.set_params(**api.DEFAULT_PARAMS)


def test_consensus_pos_large_segments(auditor):
    """Segments beyond the shared-memory segment of the parity kernel go through global scratch."""
    rng = np.random.default_rng(3)
    auditor.set_params(**api.DEFAULT_PARAMS)
    vals, offs, poss, exp = [], [0], [], []
    for n in (8192, 8193, 20000, 70000):
        pos = 1_000_000
        v = np.concatenate([pos + rng.integers(-600, 600, n // 2), rng.integers(0, 2_000_000, n - n // 2)]).astype(np.int32)
        vals.append(v); offs.append(offs[-1] + n); poss.append(pos)
        exp.append(B.consensus_pos(v, pos))
    got = auditor.consensus_pos_batch(np.concatenate(vals), offs, poss)
    assert list(got) == exp


def test_walk_quirks_golden(auditor):
    rs, queries = util.quirk_readset()
    auditor.load_reads(rs)
    for q in queries:
        auditor.set_params(consensus_min_count=q["min_count"], consensus_interval=q["interval"], consensus_interval_range=q["range"])
        sig = auditor.window_signatures(q["kind"], q["chrom"], q["start"], q["end"])
        assert [int(x) for x in sig] == q["signatures"], q
        r = auditor.refine_windows([q["chrom"]], [q["start"]], [q["end"]], [q["pos"]], [q["kind"]])
        assert int(r[0]) == q["expected"], q
    auditor.set_params(**api.DEFAULT_PARAMS)


def test_mix_windows_golden(auditor):
    rs, z = util.load_reads("mix_reads.npz")
    auditor.load_reads(rs)
    q, exp, sig, soff = z["queries"], z["expected"], z["sig"], z["sig_off"]
    groups = {}
    for i in range(len(exp)):
        groups.setdefault(tuple(int(x) for x in q[i][5:]), []).append(i)
    for (minc, interval, rg), idx in groups.items():
        auditor.set_params(consensus_min_count=minc, consensus_interval=interval, consensus_interval_range=rg)
        qq = q[idx]
        got = auditor.refine_windows(qq[:, 1], qq[:, 2], qq[:, 3], qq[:, 4], qq[:, 0])
        bad = np.flatnonzero(got != exp[idx])
        assert bad.size == 0, (minc, interval, rg, [(idx[b], int(got[b]), int(exp[idx[b]])) for b in bad[:5]])
    # raw signature multisets for a slice of the windows (the walker alone)
    for i in range(0, len(exp), 9):
        kind, chrom, st, en = (int(x) for x in q[i][:4])
        s = auditor.window_signatures(kind, chrom, st, en)
        assert np.array_equal(s, sig[soff[i]:soff[i + 1]]), i
    auditor.set_params(**api.DEFAULT_PARAMS)


def test_sliding_window_golden(auditor):
    rs, z = util.load_reads("mix_reads.npz")
    auditor.load_reads(rs)
    lines = util.load_json("mix_sliding_lines.json")
    for i, row in enumerate(z["sliding"]):
        chrom, st, en, minc, win, slide, exp = (int(x) for x in row)
        auditor.set_params(consensus_min_count=minc)
        r, ln = auditor.sliding_window_ins(chrom, st, en, win, slide, want_lines=True)
        assert r == exp, (i, r, exp)
        assert ln == lines[i], i
    auditor.set_params(**api.DEFAULT_PARAMS)


def test_process_vcf_stdout_golden(auditor):
    """Whole path: VCF text in, the reference's stdout out -- against `svtrek audt -t 1` captures."""
    rs, _ = util.load_reads("c1mini_reads.npz")
    auditor.load_reads(rs)
    gold = util.load_json("c1mini_cli.json")
    keymap = dict(interval="consensus_interval", min_count="consensus_min_count", wider="wider_interval", median="median_interval",
                  narrow="narrow_interval", range="consensus_interval_range")
    with tempfile.TemporaryDirectory() as d:
        for key, g in gold.items():
            if key == "_vcf":
                continue
            vname = key.split("/")[0]
            vcf = os.path.join(d, vname + ".vcf"); out = os.path.join(d, "out.txt")
            S.write_vcf(vcf, gold["_vcf"][vname])
            p = dict(api.DEFAULT_PARAMS); p.update({keymap[k]: v for k, v in g["params"].items()})
            auditor.set_params(**p)
            n = auditor.process_vcf(vcf, out)
            text = open(out).read()
            assert text == g["stdout"], (key, _first_diff(text, g["stdout"]))
            assert n == g["stdout"].count("\n") - 2
    auditor.set_params(**api.DEFAULT_PARAMS)


def test_process_vcf_on_a_vcf_from_the_references_simvcf_py(auditor):
    """The VCF utils/simvcf.py itself wrote (tests/golden/simvcf_cli.json) through the whole GPU path == `svtrek audt -t 1`."""
    rs, _ = util.load_reads("c1mini_reads.npz")
    auditor.load_reads(rs)
    auditor.set_params(**api.DEFAULT_PARAMS)
    gold = util.load_json("simvcf_cli.json")
    with tempfile.TemporaryDirectory() as d:
        for style, g in gold.items():
            vcf, out = os.path.join(d, style + ".vcf"), os.path.join(d, "out.txt")
            open(vcf, "w").write(g["vcf"])
            n = auditor.process_vcf(vcf, out)
            text = open(out).read()
            assert text == g["stdout"], (style, _first_diff(text, g["stdout"]))
            assert n == g["stdout"].count("\n") - 2


def _first_diff(a, b):
    for i, (x, y) in enumerate(zip(a.splitlines(), b.splitlines())):
        if x != y:
            return i, x, y
    return "length", len(a.splitlines()), len(b.splitlines())


def test_per_record_seam(auditor):
    """deletion()/insertion()/inversion() (refinement.h:25,42,60) one record at a time vs the oracle."""
    rs, _ = util.load_reads("c1mini_reads.npz")
    auditor.load_reads(rs)
    auditor.set_params(**api.DEFAULT_PARAMS)
    orc = B.OracleReads(rs)
    gold = util.load_json("c1mini_cli.json")
    n = 0
    for ln in gold["_vcf"]["endfirst"][:24]:
        f = ln.split("\t")
        pos = int(f[1])
        info = dict(kv.split("=") for kv in f[7].split(";"))
        if info["SVTYPE"] == "DEL":
            end = int(info["END"])
            got = auditor.deletion(1, (pos - 20000, pos + 2000), (end - 2000, end + 2000), (pos, end))
            want = (B.refine(orc, 0, 1, pos - 20000, pos + 2000, pos) & 0xFFFFFFFF, B.refine(orc, 1, 1, end - 2000, end + 2000, end) & 0xFFFFFFFF)
            assert got == want
            assert auditor.inversion(1, (pos - 20000, pos + 20000), (end - 20000, end + 20000), (pos, end)) == (0xFFFFFFFF, 0xFFFFFFFF)
        else:
            got = auditor.insertion(1, (pos - 10000, pos + 10000), pos)
            assert got == B.refine(orc, 2, 1, pos - 10000, pos + 10000, pos) & 0xFFFFFFFF
        n += 1
    assert n == 24
