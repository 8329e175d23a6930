#!/usr/bin/env python
"""Throughput probe of svtk_load_bam: a c2-like read set (15 kb HiFi-like reads, 30x, random SEQ/QUAL) written as BAM + BAI by
svtrek_b200.simreads.write_bam_fast, loaded under a sweep of loader settings (lanes per inflate warp, pipelined or not, copier
threads).  One JSON line per (scale, setting).  Settings are environment variables the loader reads at call time."""
import argparse
import json
import os
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scales", default="0.047,0.25,1.0", help="fractions of BASELINE config c2 (64 Mb contig, 128 000 reads)")
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--dir", default=None, help="keep / reuse the generated BAMs in this directory")
    ap.add_argument("--lanes", default="0,1,2,4,8,16,32", help="SVTK_BZ_LANES values (0: the loader's own choice)")
    ap.add_argument("--no-unpiped", action="store_true")
    ap.add_argument("--settings-json", default=None, help='instead of the built-in sweep: a JSON list of {"ENV": "value"} dicts')
    args = ap.parse_args()
    from svtrek_b200 import api, simreads as S
    d = args.dir or tempfile.mkdtemp(prefix="svtk_bam_")
    os.makedirs(d, exist_ok=True)
    a = api.Auditor(0)
    for sc in [float(x) for x in args.scales.split(",")]:
        bam = os.path.join(d, f"s{sc}.bam")
        t0 = time.perf_counter()
        rs, _ = S.make_config("c2", scale=sc)
        t1 = time.perf_counter()
        if not os.path.exists(bam + ".bai"):
            S.write_bam_fast(bam, rs)
        t2 = time.perf_counter()
        checked = False
        settings = [dict(SVTK_BZ_LANES=str(l)) for l in args.lanes.split(",")]
        if not args.no_unpiped:
            settings += [dict(SVTK_BAM_PIPE="0"), dict(SVTK_BAM_PIPE="1", SVTK_BAM_PIPE_MB="128"), dict(SVTK_BAM_PIPE="1", SVTK_BAM_PIPE_MB="256"),
                         dict(SVTK_BAM_PIPE="1", SVTK_BAM_PIPE_MB="640"), dict(SVTK_BAM_PIPE="1", SVTK_BAM_PIPE_MB="256", SVTK_BZ_LANES="8"),
                         dict(SVTK_BAM_PIPE="1", SVTK_BAM_PIPE_MB="256", SVTK_BAM_THREADS="16")]
        if args.settings_json:
            settings = json.loads(args.settings_json)
        for env in settings:
            for k in ("SVTK_BZ_LANES", "SVTK_BAM_PIPE", "SVTK_BAM_THREADS", "SVTK_BAM_PIPE_MB", "SVTK_BAM_PIPE_FIRST_MB", "SVTK_BZ_WARPS"):
                os.environ.pop(k, None)
            for k, v in env.items():
                if v != "0" or k != "SVTK_BZ_LANES":
                    os.environ[k] = v
            runs = []
            for _ in range(args.reps):
                t = time.perf_counter()
                st = a.load_bam(bam)
                st["wall_ms"] = (time.perf_counter() - t) * 1e3
                runs.append(st)
            if not checked:
                got = a.copy_reads()
                ok = all(np.array_equal(got[k].astype(np.int64), np.asarray(w).astype(np.int64)) for k, w in
                         (("pos", rs.pos), ("end", rs.end), ("n_cigar", rs.n_cigar), ("cigar", rs.cigar), ("tid_off", rs.tid_off)))
                assert ok, "loaded reads differ from the written ones"
                checked = True
            best = min(runs, key=lambda r: r["wall_ms"])
            print(json.dumps({"scale": sc, "env": env, "bam_MB": best["file_bytes"] / 1e6, "inflated_MB": best["inflated_bytes"] / 1e6, "blocks": best["n_blocks"],
                              "lanes": best["lanes_per_warp"], "launches": best["inflate_launches"], "wall_ms": best["wall_ms"], "upload_ms": best["upload_ms"],
                              "inflate_ms": best["inflate_ms"], "span_ms": best["load_span_ms"], "parse_ms": best["parse_ms"],
                              "inflate_GBps_out": best["inflated_bytes"] / best["inflate_ms"] / 1e6, "load_GBps_in": best["file_bytes"] / best["wall_ms"] / 1e6,
                              "gen_s": t1 - t0, "write_s": t2 - t1, "reads_identical": checked}), flush=True)
    a.close()


if __name__ == "__main__":
    main()
