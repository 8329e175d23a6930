#!/usr/bin/env python
"""K3 throughput probe: svtk_poa_consensus on a C4-like batch (INS 50 bp - 10 kb, 60x ONT-like), GCUPS from the kernel's
own cell counter and CUDA-event time; optional CPU abPOA timing on a sample (oracle/_ref, test infrastructure)."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from svtrek_b200 import api, simpoa as P   # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sets", type=int, default=2000)
    ap.add_argument("--len-hi", type=int, default=10000)
    ap.add_argument("--len-lo", type=int, default=50)
    ap.add_argument("--depth", type=int, default=60)
    ap.add_argument("--err", type=float, default=0.08)
    ap.add_argument("--cpu-sets", type=int, default=0)
    ap.add_argument("--reps", type=int, default=2)
    a = ap.parse_args()
    t0 = time.time()
    sets = P.ins_workload(20261107, a.sets, depth=a.depth, err=a.err, len_lo=a.len_lo, len_hi=a.len_hi)
    so, qo, b = P.pack(sets)
    print(f"generated {a.sets} sets, {len(qo)-1} reads, {len(b)/1e6:.1f} Mbases in {time.time()-t0:.1f}s", file=sys.stderr)
    out = {}
    with api.Auditor(0) as A:
        for rep in range(a.reps):
            t0 = time.time()
            cons, st = A.poa_consensus(so, qo, b)
            wall = time.time() - t0
            out = dict(st, wall_s=wall, gcups=st["cells"] / st["kernel_ms"] / 1e6, gcups_full=st["cells_full"] / st["kernel_ms"] / 1e6,
                       sets_per_s=a.sets / (st["kernel_ms"] / 1e3))
            print(json.dumps(out), file=sys.stderr)
    if a.cpu_sets:
        from oracle import binding as B
        idx = np.linspace(0, a.sets - 1, a.cpu_sets).astype(int)
        t0 = time.time()
        cells = 0
        for i in idx:
            r = B.ref_poa_steps(sets[i])
            cells += int(r["cells"].sum())
            assert np.array_equal(r["cons"], cons[i]), i
        dt = time.time() - t0
        out["cpu"] = dict(sets=len(idx), seconds=dt, gcups=cells / dt / 1e9, sets_per_s=len(idx) / dt, kind="abPOA 1.5.4 (oracle/_ref), 1 thread")
    print(json.dumps(out))


if __name__ == "__main__":
    main()
