"""SURVEY 8f-2 pinned to the reference's own BAM reader.  tests/golden/ins_sequences.npz holds, for a small seeded BAM, the
inserted sequences the REFERENCE'S HTSlib yields behind refine_ins' signatures (oracle/ins_harness.c: sam_itr_queryi +
sam_itr_next + the refine_ins loop + bam_seqi; written by tools/make_ins_golden.py).  The BAM is rebuilt here from the same seeds.

  not gpu   where /root/reference was compiled (oracle/_ref/libins_ref.so): the harness run again on the rebuilt BAM == the file
            (drift check: generator, writer and golden still agree); the plain-Python checker tests/util_bam.py == the file too
  gpu       svtk_ins_sequences through the C ABI == the file, for three (tol, min_len) settings"""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
from make_ins_golden import ins_case  # noqa: E402

from oracle import binding as B  # noqa: E402
from tests import util  # noqa: E402

SETTINGS = ((5, 50), (0, 50), (40, 120))


def golden(tol, min_len):
    z = np.load(os.path.join(util.GOLD, "ins_sequences.npz"))
    return {k: z[f"t{tol}_m{min_len}_{k}"] for k in ("window", "read_pos", "ref_pos", "seq_off", "bases")}, z["refined"]


def test_golden_is_what_the_reference_htslib_yields_and_the_python_checker_agrees(tmp_path):
    from tests import util_bam as U
    bam = str(tmp_path / "ins.bam")
    rs, truth, w, refined = ins_case(bam)
    _, gref = golden(5, 50)
    assert np.array_equal(refined, gref), "the oracle's refined positions drifted from the golden file's"
    rd = U.read_bam(bam)
    for tol, min_len in SETTINGS:
        g, _ = golden(tol, min_len)
        if B.have_ins_ref():
            r = B.ref_ins_sequences(bam, w["chrom"], w["start"], w["end"], w["kind"], refined, tol=tol, min_len=min_len)
            for k in g:
                assert np.array_equal(r[k], g[k]), (tol, min_len, k)
        flat = [(win, int(rd["pos"][r]), rp, codes) for win, seqs in U.ins_sequences(rd, w["chrom"], w["start"], w["end"], w["kind"], refined, tol=tol, min_len=min_len)
                for r, rp, codes in seqs]
        assert len(flat) == len(g["window"])
        for k, (win, rpos, rp, codes) in enumerate(flat):
            assert (win, rpos, rp) == (int(g["window"][k]), int(g["read_pos"][k]), int(g["ref_pos"][k]))
            assert np.array_equal(codes, g["bases"][int(g["seq_off"][k]):int(g["seq_off"][k + 1])])


@pytest.mark.gpu
def test_svtk_ins_sequences_equal_the_reference_htslib(tmp_path):
    from svtrek_b200 import api
    bam = str(tmp_path / "ins.bam")
    rs, truth, w, refined = ins_case(bam)
    with api.Auditor(0) as a:
        a.load_bam(bam)
        got_ref = a.refine_windows(w["chrom"], w["start"], w["end"], w["pos"], w["kind"])
        assert np.array_equal(got_ref, refined)
        for tol, min_len in SETTINGS:
            g, _ = golden(tol, min_len)
            got = a.ins_sequences(w["chrom"], w["start"], w["end"], w["pos"], w["kind"], refined, tol=tol, min_len=min_len)
            assert len(got["read"]) == len(g["window"]) and len(g["window"]) > 100
            win_of_seq = np.repeat(got["window"], np.diff(got["set_off"]))
            assert np.array_equal(win_of_seq, g["window"])
            assert np.array_equal(rs.pos[got["read"]], g["read_pos"])
            assert np.array_equal(got["ref_pos"], g["ref_pos"])
            assert np.array_equal(got["seq_off"], g["seq_off"]) and np.array_equal(got["bases"], g["bases"])
