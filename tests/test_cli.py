"""svtrek_gpu (csrc/svtrek_gpu.c, the C host driver over the C ABI) against the UNMODIFIED reference binary, live:
same BAM, same VCF, same options -> the same stdout, byte for byte.  The reference runs with -t 1 on a VCF padded past its
end-of-file tail drop, which is the output the GPU driver reproduces for every record."""
import os
import subprocess

import pytest

from svtrek_b200 import simreads as S

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CLI = os.path.join(ROOT, "svtrek_b200", "svtrek_gpu")


def test_cli_is_built_and_refuses_to_run_without_its_inputs():
    from svtrek_b200 import build
    build.build()
    assert os.path.exists(CLI)
    p = subprocess.run([CLI, "audt", "-h"], stdout=subprocess.PIPE, stderr=subprocess.PIPE)
    assert p.returncode == 0 and b"--consensus-min-count" in p.stdout and b"--wider-interval" in p.stdout
    p = subprocess.run([CLI, "audt", "-v", "/dev/null"], stdout=subprocess.PIPE, stderr=subprocess.PIPE)
    assert p.returncode != 0 and b"[ERROR] BAM file is not provided." in p.stderr
    p = subprocess.run([CLI, "disc"], stdout=subprocess.PIPE, stderr=subprocess.PIPE)
    assert p.returncode != 0


@pytest.mark.gpu
@pytest.mark.parametrize("extra", [[], ["--consensus-min-count", "2", "--narrow-interval", "1000", "--consensus-interval", "10"],
                                   ["--wider-interval", "5000", "--median-interval", "3000", "--consensus-interval-range", "100"]])
def test_cli_stdout_equals_the_reference_binary(tmp_path, extra):
    from oracle import binding as B
    if not (B.have_ref() and os.path.exists(os.path.join(B.REF_DIR, "svtrek"))):
        pytest.skip("oracle/_ref/svtrek not built here")
    rs, truth = S.make_config("c1", scale=0.25)
    bam, vcf = str(tmp_path / "s.bam"), str(tmp_path / "s.vcf")
    S.write_bam(bam, rs, truth=truth)
    S.write_vcf(vcf, S.vcf_lines(truth, end_first=bool(extra)), pad=16)
    ref_out, _ = B.run_ref_cli(bam, vcf, threads=1, extra=extra)
    p = subprocess.run([CLI, "audt", "-b", bam, "-v", vcf, "-t", "1", *extra], stdout=subprocess.PIPE, stderr=subprocess.PIPE, check=True)
    assert p.stdout.decode() == ref_out
    if not extra:                                            # -o: the same lines in the file as well
        o = str(tmp_path / "lines.out")
        p2 = subprocess.run([CLI, "audt", "-b", bam, "-v", vcf, "-o", o], stdout=subprocess.PIPE, stderr=subprocess.PIPE, check=True)
        assert p2.stdout.decode() == ref_out and open(o).read() == ref_out
    n = ref_out.count("(DEL) ") + ref_out.count("(INS) ")
    assert n >= 100 and ref_out.startswith("[INFO] Started") and "NA" in ref_out or n >= 100


@pytest.mark.gpu
def test_cli_ins_consensus_side_file(tmp_path):
    """--ins-consensus: stdout untouched; one FASTA entry per refined INS record with carriers, equal to svtk_ins_sequences +
    svtk_poa_consensus through the API, and close to the planted insertion (same length within 2 %)."""
    import numpy as np
    from svtrek_b200 import api
    rs, truth = S.make_config("c1", scale=0.25)
    bam, vcf, fa = str(tmp_path / "s.bam"), str(tmp_path / "s.vcf"), str(tmp_path / "ins.fa")
    S.write_bam(bam, rs, truth=truth)
    S.write_vcf(vcf, S.vcf_lines(truth), pad=0)
    plain = subprocess.run([CLI, "audt", "-b", bam, "-v", vcf], stdout=subprocess.PIPE, stderr=subprocess.PIPE, check=True).stdout
    p = subprocess.run([CLI, "audt", "-b", bam, "-v", vcf, "--ins-consensus", fa, "--stats"], stdout=subprocess.PIPE, stderr=subprocess.PIPE, check=True)
    assert p.stdout == plain and b"insertion consensus sequences" in p.stderr
    entries = open(fa).read().strip().split("\n")
    heads, seqs = entries[0::2], entries[1::2]
    assert len(heads) == len(seqs) > 20 and all(h.startswith(">INS chr=1 org_pos=") for h in heads)
    # the same through the API: windows exactly as process_vcf builds them for INS records (pos +- median_interval)
    a = api.Auditor(0)
    a.load_bam(bam)
    by_pos = {}
    for h, q in zip(heads, seqs):
        f = dict(x.split("=") for x in h[5:].split())
        assert int(f["len"]) == len(q)
        by_pos[int(f["org_pos"])] = (int(f["ref_pos"]), int(f["reads"]), q)
    pos = np.array(sorted(by_pos), np.int64)
    n = len(pos)
    w = dict(chrom=np.ones(n, np.int32), start=pos - 10000, end=pos + 10000, pos=pos, kind=np.full(n, 2, np.uint8))
    refined = a.refine_windows(w["chrom"], w["start"], w["end"], w["pos"], w["kind"])
    sets = a.ins_sequences(w["chrom"], w["start"], w["end"], w["pos"], w["kind"], refined, tol=5)
    cons, _ = a.poa_consensus(sets["set_off"], sets["seq_off"], sets["bases"])
    assert len(cons) == n
    for s, wi in enumerate(sets["window"]):
        ref_pos, reads, q = by_pos[int(pos[wi])]
        assert ref_pos == int(refined[wi]) and reads == int(sets["set_off"][s + 1] - sets["set_off"][s])
        assert q == "".join("ACGTN"[c] for c in cons[s])
    ins_len = {int(p0): int(L) for p0, L, d in zip(truth.pos0, truth.length, truth.is_del) if not d}
    assert sum(1 for q in seqs if any(abs(len(q) - L) <= max(2, L // 50) for L in ins_len.values())) >= len(seqs) - 2
