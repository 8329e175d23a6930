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
