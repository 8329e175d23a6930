#!/usr/bin/env python
"""Throughput probe of svtk_load_bam: a c2-like read set (15 kb HiFi-like reads, 30x, random SEQ/QUAL) written as BAM by the
reference's HTSlib tools (oracle/_ref), loaded `--reps` times; zlib's single-core inflate of the same blocks beside it."""
import argparse
import json
import os
import sys
import tempfile
import time
import zlib

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scale", type=float, default=0.125, help="fraction of BASELINE config c2 (64 Mb contig, 128 000 reads)")
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--cpu-blocks", type=int, default=400)
    ap.add_argument("--dir", default=None, help="keep / reuse the generated BAM in this directory")
    args = ap.parse_args()
    from oracle import binding as B
    from svtrek_b200 import api, simreads as S
    from tests import util_bam as U
    rs, _ = S.make_config("c2", scale=args.scale)
    d = args.dir or tempfile.mkdtemp(prefix="svtk_bam_")
    os.makedirs(d, exist_ok=True)
    sam, bam = os.path.join(d, "s.sam"), os.path.join(d, f"s{args.scale}.bam")
    t0 = t1 = t2 = time.perf_counter()
    if not os.path.exists(bam + ".bai"):
        S.write_sam(sam, rs); t1 = time.perf_counter(); B.make_bam(sam, bam); t2 = time.perf_counter()
        os.remove(sam)
    a = api.Auditor(0)
    runs = []
    for _ in range(args.reps):
        t = time.perf_counter()
        st = a.load_bam(bam)
        st["wall_ms"] = (time.perf_counter() - t) * 1e3
        runs.append(st)
    got = a.copy_reads()
    ok = all(np.array_equal(got[k].astype(np.int64), np.asarray(w).astype(np.int64)) for k, w in
             (("pos", rs.pos), ("end", rs.end), ("n_cigar", rs.n_cigar), ("cigar", rs.cigar), ("tid_off", rs.tid_off)))
    buf = open(bam, "rb").read()
    blocks = U.bgzf_blocks(buf)[:args.cpu_blocks]
    t = time.perf_counter()
    nb = sum(len(zlib.decompress(buf[p:p + n], -15)) for _, p, n, _ in blocks)
    cpu_s = time.perf_counter() - t
    best = min(runs, key=lambda r: r["total_ms"])
    print(json.dumps({"reads": rs.n_reads, "bam_MB": len(buf) / 1e6, "inflated_MB": best["inflated_bytes"] / 1e6, "blocks": best["n_blocks"],
                      "pieces": best["n_pieces"], "identical_to_written_reads": bool(ok), "runs": runs,
                      "inflate_GBps_out": best["inflated_bytes"] / best["inflate_ms"] / 1e6, "total_GBps_out": best["inflated_bytes"] / best["total_ms"] / 1e6,
                      "zlib_1core_GBps_out": nb / cpu_s / 1e9, "make_sam_s": t1 - t0, "make_bam_s": t2 - t1}))


if __name__ == "__main__":
    main()
