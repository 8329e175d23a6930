"""Host logic of the batch seam on CPU: the VCF front end (svtk_vcf_windows: thread_func's parsing + window construction,
audit.c:65-193) and the stdout back end (svtk_vcf_format: audit.c:182-229) of libsvtk_b200.so, with the CPU oracle standing in
for the GPU between them, against the stdout / stderr the unmodified reference printed (`svtrek audt -t 1`, tests/golden).
No device is touched: these two entry points take no ctx."""
import os

import numpy as np
import pytest

from oracle import binding as B
from svtrek_b200 import api
from svtrek_b200 import simreads as S
from tests import util

KEYMAP = dict(interval="consensus_interval", min_count="consensus_min_count", wider="wider_interval", median="median_interval",
              narrow="narrow_interval", range="consensus_interval_range")


def _audit_on_cpu(vcf, out, orc, params, n_ref, refs=None, owns_unplaced=True, info_lines=True, capfd=None):
    w = api.vcf_windows(vcf, params, n_ref=n_ref, refs=refs, owns_unplaced=owns_unplaced)
    if capfd is not None:
        capfd.readouterr()                     # every parse of the file reports its malformed lines: keep the last one's
    res = B.refine_batch(orc, w["chrom"], w["start"], w["end"], w["pos"], w["kind"], min_count=params["consensus_min_count"],
                         interval=params["consensus_interval"], range_=params["consensus_interval_range"])
    return api.vcf_format(vcf, res, out, params, n_ref=n_ref, refs=refs, owns_unplaced=owns_unplaced, info_lines=info_lines), w


@pytest.mark.parametrize("threads", ["1", "5"])
def test_front_and_back_end_equal_the_reference_cli(lib_built, tmp_path, capfd, monkeypatch, threads):
    monkeypatch.setenv("SVTK_VCF_THREADS", threads)
    rs, _ = util.load_reads("c1mini_reads.npz")
    orc = B.OracleReads(rs)
    gold = util.load_json("c1mini_cli.json")
    for key, g in gold.items():
        if key == "_vcf":
            continue
        vname = key.split("/")[0]
        vcf, out = str(tmp_path / (vname + ".vcf")), str(tmp_path / "out.txt")
        S.write_vcf(vcf, gold["_vcf"][vname])
        p = dict(api.DEFAULT_PARAMS); p.update({KEYMAP[k]: v for k, v in g["params"].items()})
        n, _ = _audit_on_cpu(vcf, out, orc, p, rs.n_tid, capfd=capfd)
        err = capfd.readouterr().err
        text = open(out).read()
        assert text == g["stdout"], key
        assert n == g["stdout"].count("\n") - 2
        # the reference's stderr for the same lines, in order
        assert err == g["stderr"], key


def test_region_shards_concatenate_to_the_whole(lib_built, tmp_path, monkeypatch):
    """svtk_process_vcf_region's rule: a record belongs to the shard holding its reference; what no shard can place goes to the
    owner of the unplaced.  Three shards + merge == the unsharded output, byte for byte."""
    rs, truth = S.simulate(11, [400_000, 300_000, 350_000, 200_000], 20, 15000, 0.005, 160, margin=30000)
    orc = B.OracleReads(rs)
    lines = S.vcf_lines(truth, end_first=True)
    lines.insert(5, "chrX\t5000\t.\tN\tN\t.\tPASS\tSVTYPE=INS")                        # chrom 0: no shard holds it
    lines.insert(9, "chr9\t70000\t.\tN\tN\t.\tPASS\tEND=70400;SVTYPE=DEL")            # beyond the dictionary
    lines.insert(12, "junk")                                                          # "VCF: no index"
    lines.insert(20, "chr2\t100\t.\tN\tN\t.\tPASS\tSVTYPE=DUP")                       # "[ERROR] Unkown type." (owned by chr2's shard)
    vcf = str(tmp_path / "j.vcf")
    S.write_vcf(vcf, lines)
    p = dict(api.DEFAULT_PARAMS)
    whole = str(tmp_path / "whole.txt")
    n_all, w_all = _audit_on_cpu(vcf, whole, orc, p, rs.n_tid)
    want = open(whole).read()
    for threads in ("1", "4"):
        monkeypatch.setenv("SVTK_VCF_THREADS", threads)
        parts, n_sum, n_win = [], 0, 0
        for k, refs in enumerate([(0, 0), (1, 2), (3, 3)]):
            o = str(tmp_path / f"part{k}.txt")
            n, w = _audit_on_cpu(vcf, o, orc, p, rs.n_tid, refs=refs, owns_unplaced=(k == 0), info_lines=False)
            parts.append(open(o).read()); n_sum += n; n_win += len(w["chrom"])
        assert n_sum == n_all and n_win == len(w_all["chrom"])
        body = want.split("\n", 1)[1].rsplit("[INFO] Ended", 1)[0]
        # shard order is reference order; the VCF is sorted by reference apart from the planted oddities, which the first shard owns
        assert sorted("".join(parts).splitlines()) == sorted(body.splitlines())
        merged = "".join(parts)
        placed = [ln for ln in body.splitlines() if not (ln.startswith("(INS) chr: 0") or ln.startswith("(DEL) chr: 9"))]
        assert [ln for ln in merged.splitlines() if not (ln.startswith("(INS) chr: 0") or ln.startswith("(DEL) chr: 9"))] == placed


def test_many_pieces_equal_one_piece(lib_built, tmp_path, monkeypatch):
    """The parallel line split: 16 pieces cut at arbitrary byte offsets give the windows and the lines of a serial parse, incl.
    a last line without a newline, empty lines, comment lines and CRLF-free junk."""
    rs, truth = S.simulate(3, [900_000], 10, 15000, 0.005, 500, margin=30000, min_gap=200)
    lines = S.vcf_lines(truth, end_first=False)
    lines[7:7] = ["", "#comment in the middle", "x", "chr1\tabc\t.\tN\tN\t.\tPASS\tSVTYPE=INS"]
    vcf = str(tmp_path / "m.vcf")
    with open(vcf, "w") as f:
        f.write(S.VCF_HEADER + "\n".join(lines))                                      # no trailing newline
    res = {}
    for threads in ("1", "16"):
        monkeypatch.setenv("SVTK_VCF_THREADS", threads)
        w = api.vcf_windows(vcf)
        out = str(tmp_path / f"o{threads}.txt")
        api.vcf_format(vcf, np.arange(len(w["chrom"]), dtype=np.int32) % 7 - 1, out)
        res[threads] = (w, open(out).read())
    for k in ("chrom", "start", "end", "pos", "kind"):
        assert np.array_equal(res["1"][0][k], res["16"][0][k]), k
    assert res["1"][1] == res["16"][1]
    assert len(res["1"][0]["chrom"]) > 400


def test_a_vcf_written_by_the_references_own_simvcf_py(lib_built, tmp_path, capfd):
    """tests/golden/simvcf_cli.json: utils/simvcf.py (unmodified, tools/make_simvcf_golden.py) wrote the VCF -- header lines, CIPOS /
    CIEND before END (END parses as CIEND's value), 'chr' stripped from CHROM, untouched short variants -- and
    `svtrek audt -t 1` printed the stdout.  Front end + oracle + back end reproduce it byte for byte."""
    rs, _ = util.load_reads("c1mini_reads.npz")
    orc = B.OracleReads(rs)
    gold = util.load_json("simvcf_cli.json")
    for style, g in gold.items():
        vcf, out = str(tmp_path / (style + ".vcf")), str(tmp_path / "o.txt")
        open(vcf, "w").write(g["vcf"])
        n, w = _audit_on_cpu(vcf, out, orc, dict(api.DEFAULT_PARAMS), rs.n_tid, capfd=capfd)
        err = capfd.readouterr().err
        assert open(out).read() == g["stdout"], style
        assert err == g["stderr"], style
        assert n == g["stdout"].count("\n") - 2 and n >= 90
