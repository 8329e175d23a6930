"""Region shards of one job, file to file (svtk_audit_shard / svtk_load_bam_windows / svtrek_gpu --devices): the multi-GPU form of
process_vcf (audit.c:250; worker fan-out audit.c:269-293).  A shard loads only what its own windows can see -- looked up in the
.bai like hts_itr_query does (hts.c:3049-3090) -- so cuts may fall inside a reference.  Checked here on ONE device (the shards run
one after the other on the same ctx): shard outputs concatenated == the unsharded output == the oracle's."""
import os
import subprocess

import numpy as np
import pytest

from oracle import binding as B
from svtrek_b200 import api
from svtrek_b200 import simreads as S

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def job(tmp_path_factory):
    d = tmp_path_factory.mktemp("shardjob")
    # two references, one of them long enough that 5 shards must cut inside it
    rs, truth = S.simulate(77, [1_400_000, 300_000], 20, 12000, 0.005, 260, margin=30000)
    bam, vcf = str(d / "j.bam"), str(d / "j.vcf")
    S.write_bam_fast(bam, rs, truth=truth)
    S.write_vcf(vcf, S.vcf_lines(truth, end_first=True), pad=8)
    return rs, truth, bam, vcf, d


def test_windows_load_is_a_superset_that_refines_like_the_whole(job):
    rs, truth, bam, vcf, d = job
    w = S.windows_from_truth(truth)
    want = B.refine_batch(B.OracleReads(rs), w["chrom"], w["start"], w["end"], w["pos"], w["kind"])
    with api.Auditor(0) as a:
        whole = a.load_bam(bam)
        assert np.array_equal(a.refine_windows(w["chrom"], w["start"], w["end"], w["pos"], w["kind"]), want)
        # a third of the windows, from the middle of reference 1: far fewer bytes shipped, same results
        lo, hi = len(w["chrom"]) // 3, 2 * len(w["chrom"]) // 3
        sel = slice(lo, hi)
        st = a.load_bam_windows(bam, w["chrom"][sel], w["start"][sel], w["end"][sel])
        assert 0 < st["image_bytes"] < 0.7 * whole["image_bytes"] and st["n_reads"] < whole["n_reads"]
        got = a.refine_windows(w["chrom"][sel], w["start"][sel], w["end"][sel], w["pos"][sel], w["kind"][sel])
        assert np.array_equal(got, want[sel])
        # every loaded read is a read of the file, in file order
        rd = a.copy_reads()
        assert (np.diff(rd["pos"][rd["tid_off"][0]:rd["tid_off"][1]].astype(np.int64)) >= 0).all()
        # windows nobody can serve (wrapped coordinates, unknown reference): nothing to load, NA
        st0 = a.load_bam_windows(bam, [9, 1], [5, 4294967000], [9, 200])
        assert st0["n_reads"] == 0
        assert list(a.refine_windows([1], [100000], [130000], [120000], [0])) == [-1]


@pytest.mark.parametrize("n_shards", [1, 2, 5])
def test_shards_concatenate_to_the_whole_output(job, n_shards):
    rs, truth, bam, vcf, d = job
    whole = str(d / "whole.txt")
    with api.Auditor(0) as a:
        a.load_bam(bam)
        n_all = a.process_vcf(vcf, whole)
        text = open(whole).read()
        parts, n_sum, shipped = [], 0, []
        for k in range(n_shards):
            o = str(d / f"shard{k}.txt")
            n, st = a.audit_shard(bam, vcf, o, k, n_shards)
            parts.append(open(o).read()); n_sum += n; shipped.append(st["image_bytes"])
    assert n_sum == n_all
    assert "[INFO] Started processing variation file.\n" + "".join(parts) + "[INFO] Ended processing variation file\n" == text
    if n_shards == 5:
        assert max(shipped) < 0.5 * os.path.getsize(bam)       # the cuts really fall inside reference 1
        assert min(len(p) for p in parts) > 0


def test_cli_devices_mode_equals_single_device(job):
    rs, truth, bam, vcf, d = job
    cli = os.path.join(ROOT, "svtrek_b200", "svtrek_gpu")
    one = subprocess.run([cli, "audt", "-b", bam, "-v", vcf], capture_output=True, check=True)
    out = str(d / "tee.txt")
    many = subprocess.run([cli, "audt", "-b", bam, "-v", vcf, "--devices", "0,0,0", "-o", out], capture_output=True, check=True)
    assert many.stdout == one.stdout and len(one.stdout) > 1000
    assert open(out, "rb").read() == one.stdout              # -o: an additional sink
    assert many.stderr == one.stderr                         # the pad lines' messages once, from shard 0
    tee = subprocess.run([cli, "audt", "-b", bam, "-v", vcf, "-o", out], capture_output=True, check=True)
    assert tee.stdout == one.stdout and open(out, "rb").read() == one.stdout
