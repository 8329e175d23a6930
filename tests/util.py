"""Shared helpers for the test-suite: golden-vector loading."""
import json
import os

import numpy as np

from svtrek_b200.simreads import ReadSet

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_reads(name):
    z = np.load(os.path.join(GOLD, name))
    rs = ReadSet(z["tid_off"], z["pos"], z["end"], z["n_cigar"], z["cigar_off"], z["cigar"], [int(x) for x in z["contig_len"]])
    return rs, z


def load_json(name):
    return json.load(open(os.path.join(GOLD, name)))


def quirk_readset():
    """The single-read contigs of tests/golden/walk_quirks.json as a ReadSet (CIGAR strings parsed here)."""
    from svtrek_b200 import simreads as S
    g = load_json("walk_quirks.json")
    reads = sorted(g["reads"], key=lambda r: (r[0], r[1]))
    words, off, pos, ncig = [], [0], [], []
    tid_off = [0] * (len(g["contig_len"]) + 1)
    for t, p, c in reads:
        num = 0
        w = []
        for ch in c:
            if ch.isdigit():
                num = num * 10 + int(ch)
            else:
                w.append(num << 4 | S.OPCHARS.index(ch)); num = 0
        words += w; off.append(len(words)); pos.append(p); ncig.append(len(w)); tid_off[t + 1] += 1
    rs = ReadSet(np.cumsum(tid_off).astype(np.int64), np.array(pos, dtype=np.uint32), None, np.array(ncig, dtype=np.uint32),
                 np.array(off, dtype=np.uint64), np.array(words, dtype=np.uint32), g["contig_len"])
    rs.end = S.endpos_from_cigar(rs.pos, rs.cigar_off, rs.cigar)
    return rs, g["queries"]
