#!/usr/bin/env python
"""tests/golden/simvcf_cli.json: a VCF written by the REFERENCE'S OWN utils/simvcf.py (run unmodified from /root/reference) and
what `svtrek audt -t 1` prints for it against the c1mini read set (tests/golden/c1mini_reads.npz).

  precise truth VCF (real REF / ALT strings, three ways of naming the SV type)  --simvcf.py-->  imprecise VCF  --svtrek-->  stdout

simvcf.py draws its confidence intervals from an unseeded random(), so its OUTPUT is the fixture: the file is stored verbatim.
Run here (needs /root/reference and oracle/_ref); the GPU box only reads the .json."""
import json
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
GOLD = os.path.join(ROOT, "tests", "golden")
SIMVCF = "/root/reference/utils/simvcf.py"


def precise_vcf(truth, style):
    """style 'len': no SV type in INFO (simvcf.py compares REF / ALT lengths); 'tag': SVTYPE=DEL / SVTYPE=INS present."""
    rng = np.random.default_rng(17)
    head = ("##fileformat=VCFv4.2\n" '##INFO=<ID=SVTYPE,Number=1,Type=String,Description="Type of structural variant">\n'
            '##INFO=<ID=SVLEN,Number=1,Type=Integer,Description="Length">\n' "#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO\tFORMAT\tSAMPLE\n")
    out = [head]
    for i in range(truth.pos0.shape[0]):
        ln, p = int(truth.length[i]), int(truth.pos0[i])
        seq = "".join("ACGT"[k] for k in rng.integers(0, 4, ln + 1))
        ref, alt = (seq, seq[0]) if truth.is_del[i] else (seq[0], seq)
        info = f"SVLEN={'-' if truth.is_del[i] else ''}{ln}"
        if style == "tag":
            info = f"SVTYPE={'DEL' if truth.is_del[i] else 'INS'};" + info
        out.append(f"chr1\t{p}\t.\t{ref}\t{alt}\t.\tPASS\t{info}\tGT\t0/1\n")
    # a few records simvcf.py passes through untouched (short indel, SNV) -- svtrek skips them silently
    out.append("chr1\t1234\t.\tA\tAT\t.\tPASS\tSVLEN=1\tGT\t0/1\n")
    out.append("chr1\t2345\t.\tA\tC\t.\tPASS\t.\tGT\t0/1\n")
    return "".join(out)


def main():
    from oracle import binding as B
    from svtrek_b200 import simreads as S
    from make_golden import to_bam
    assert B.have_ref() and os.path.exists(SIMVCF)
    # the read set is the stored c1mini one; the "truth" are the calls of its stored VCF (POS / END / SVLEN as written there)
    from tests import util
    rs, _ = util.load_reads("c1mini_reads.npz")
    lines = json.load(open(os.path.join(GOLD, "c1mini_cli.json")))["_vcf"]["endfirst"]
    pos0, length, is_del = [], [], []
    for ln in lines:
        f = ln.split("\t")
        info = dict(kv.split("=") for kv in f[7].split(";"))
        pos0.append(int(f[1])); is_del.append(info["SVTYPE"] == "DEL"); length.append(abs(int(info["SVLEN"])))
    truth = S.SVTruth(np.zeros(len(pos0), np.int32), np.array(pos0), np.array(length), np.array(is_del))
    gold = {}
    with tempfile.TemporaryDirectory() as tmp:
        bam = to_bam(rs, tmp, "c1mini", random_seq=False)
        for style, extra in (("len", []), ("tag", ["--tag", "SVTYPE"])):
            src, dst = os.path.join(tmp, style + ".vcf"), os.path.join(tmp, style + ".sim.vcf")
            open(src, "w").write(precise_vcf(truth, style))
            r = subprocess.run([sys.executable, SIMVCF, "-i", src, "-o", dst] + extra, capture_output=True, text=True, cwd=tmp)
            assert r.returncode == 0 and os.path.exists(dst), r.stdout + r.stderr
            text = open(dst).read()
            padded = os.path.join(tmp, style + ".pad.vcf")
            open(padded, "w").write(text + "pad\n" * 16)            # the reference drops up to 2t queued lines at end of file
            out, err = B.run_ref_cli(bam, padded, threads=1)
            err = "".join(ln + "\n" for ln in err.splitlines() if ln != "VCF: no index at line: pad")
            n = out.count("(DEL)") + out.count("(INS)")
            assert n > 80 and "org end: 42949" in out, "simvcf.py's CIEND-before-END layout should make END parse as CIEND's value (SURVEY F4)"
            gold[style] = dict(vcf=text, stdout=out, stderr=err, simvcf_args=extra)
            print(f"{style}: simvcf.py wrote {text.count(chr(10))} lines; svtrek printed {n} records, {out.count('ref end: NA')} with ref end NA")
    json.dump(gold, open(os.path.join(GOLD, "simvcf_cli.json"), "w"), indent=0)
    print("wrote tests/golden/simvcf_cli.json")


if __name__ == "__main__":
    main()
