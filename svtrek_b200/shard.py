"""Region sharding of one audt job across GPUs (BASELINE north star: "VCF records are sharded by genomic region across
the 8 GPUs ... no NCCL collective is needed and per-GPU results are merged in order on the host").

The reference's unit of fan-out is one VCF line per pool thread (audit.c:289-338, tpool.c); here it is one contiguous
run of VCF records per GPU.  A shard is closed under its reads: it takes every read an iterator of one of its windows
could yield -- pos < max(end) - 1 and endpos > min(start) - 1 (hts.c:3913-3916), found with the same look-back bound the
device search uses (the longest reference span of any read) -- so a rank needs nothing from its neighbours and the only
exchange is the gather of 4-byte results in shard order == VCF order.
"""
from __future__ import annotations

import numpy as np

from .simreads import ReadSet


def window_work(windows: dict, rs: ReadSet = None):
    """Estimated work per window: the reads its iterator yields (width x local depth) when the read set is at hand -- a 200x tandem
    cluster weighs what it costs (config 5 is skewed) -- else its width."""
    width = (windows["end"].astype(np.int64) - windows["start"].astype(np.int64)) & 0xFFFFFFFF
    width = np.where(width > (1 << 31), 0, width) + 1          # a uint32-wrapped window sees no reads
    if rs is None or rs.n_reads == 0:
        return width.astype(np.float64)
    span = int((rs.end.astype(np.int64) - rs.pos.astype(np.int64)).max(initial=1))
    work = np.ones(width.shape[0], dtype=np.float64)
    chrom = np.asarray(windows["chrom"]).astype(np.int64)
    for t in np.unique(chrom - 1):
        if t < 0 or t >= rs.n_tid:
            continue
        m = np.flatnonzero((chrom - 1 == t) & (width > 1))
        p = rs.pos[int(rs.tid_off[t]):int(rs.tid_off[t + 1])].astype(np.int64)
        lo = np.searchsorted(p, windows["start"][m].astype(np.int64) - span, side="left")
        hi = np.searchsorted(p, windows["end"][m].astype(np.int64), side="left")
        work[m] += np.maximum(hi - lo, 0)
    return work


def plan_shards(windows: dict, n_shards: int, rs: ReadSet = None):
    """Cut the window list (VCF order) into n_shards contiguous pieces of roughly equal work (window_work), never splitting a
    record's windows (a DEL owns two).  Returns [(win_lo, win_hi)] of length n_shards (possibly empty ranges)."""
    n = int(len(windows["chrom"]))
    if n == 0:
        return [(0, 0)] * n_shards
    cum = np.cumsum(window_work(windows, rs))
    rec = np.asarray(windows["record"])
    cuts = [0]
    for s in range(1, n_shards):
        i = int(np.searchsorted(cum, cum[-1] * s / n_shards))
        i = min(max(i, cuts[-1]), n)
        while 0 < i < n and rec[i] == rec[i - 1]:               # keep a record's windows together
            i += 1
        cuts.append(i)
    cuts.append(n)
    return [(cuts[i], cuts[i + 1]) for i in range(n_shards)]


def shard_reads(rs: ReadSet, windows: dict, lo: int, hi: int) -> ReadSet:
    """The sub read set windows[lo:hi] can see (all contigs kept, so chrom numbers stay valid)."""
    span = int((rs.end.astype(np.int64) - rs.pos.astype(np.int64)).max(initial=1))
    n_tid = rs.n_tid
    keep_lo = np.zeros(n_tid, np.int64); keep_hi = np.zeros(n_tid, np.int64)
    for t in range(n_tid):
        keep_lo[t] = keep_hi[t] = rs.tid_off[t]
    chrom = np.asarray(windows["chrom"][lo:hi]).astype(np.int64)
    ws = np.asarray(windows["start"][lo:hi]).astype(np.int64)
    we = np.asarray(windows["end"][lo:hi]).astype(np.int64)
    for t in np.unique(chrom - 1):
        if t < 0 or t >= n_tid:
            continue
        m = chrom - 1 == t
        beg = ws[m] - 1; end = we[m] - 1                        # sam_itr_queryi(idx, tid, start - 1, end - 1), refinement.c:114
        ok = (end >= beg) & (beg >= 0)
        if not ok.any():
            continue
        t0, t1 = int(rs.tid_off[t]), int(rs.tid_off[t + 1])
        p = rs.pos[t0:t1].astype(np.int64)
        keep_lo[t] = t0 + int(np.searchsorted(p, beg[ok].min() - span + 1, side="left"))
        keep_hi[t] = t0 + int(np.searchsorted(p, end[ok].max(), side="left"))
        keep_hi[t] = max(keep_hi[t], keep_lo[t])
    idx = np.concatenate([np.arange(keep_lo[t], keep_hi[t]) for t in range(n_tid)]) if n_tid else np.zeros(0, np.int64)
    tid_off = np.zeros(n_tid + 1, np.int64)
    tid_off[1:] = np.cumsum(keep_hi - keep_lo)
    ncig = rs.n_cigar[idx].astype(np.uint32)
    coff = np.zeros(len(idx) + 1, np.uint64)
    coff[1:] = np.cumsum(ncig.astype(np.uint64))
    words = [rs.cigar[int(rs.cigar_off[keep_lo[t]]):int(rs.cigar_off[keep_hi[t]])] for t in range(n_tid) if keep_hi[t] > keep_lo[t]]
    cigar = np.concatenate(words).astype(np.uint32) if words else np.zeros(0, np.uint32)
    return ReadSet(tid_off, rs.pos[idx], rs.end[idx], ncig, coff, cigar, list(rs.contig_len))


def audit_sharded(refine_fn, rs: ReadSet, windows: dict, rank: int, world: int, merge_dir: str = None, group=None):
    """Run this rank's shard through refine_fn(sub_reads, chrom, start, end, pos, kind) -> int32 results; rank 0 returns the whole
    job's results in VCF order, the other ranks None.  The merge happens ON THE HOST: every rank leaves its 4-byte results in a
    file of `merge_dir` (one node: a directory all ranks see), rank 0 concatenates them in shard order == VCF order.  No
    collective touches the data path (north star); torch.distributed only provides the barrier."""
    import os
    import tempfile
    shards = plan_shards(windows, world, rs)
    lo, hi = shards[rank]
    sub = shard_reads(rs, windows, lo, hi)
    mine = np.asarray(refine_fn(sub, *[windows[k][lo:hi] for k in ("chrom", "start", "end", "pos", "kind")]), dtype=np.int32)
    if world == 1:
        return mine
    import torch.distributed as dist
    merge_dir = merge_dir or os.path.join(tempfile.gettempdir(), "svtk_merge_" + os.environ.get("MASTER_PORT", "0"))
    os.makedirs(merge_dir, exist_ok=True)
    tmp = os.path.join(merge_dir, f"shard{rank}.tmp.npy")
    np.save(tmp, mine)
    os.replace(tmp, os.path.join(merge_dir, f"shard{rank}.npy"))
    dist.barrier(group=group)
    out = None
    if rank == 0:
        out = np.concatenate([np.load(os.path.join(merge_dir, f"shard{r}.npy")) for r in range(world)])
        assert out.shape[0] == len(windows["chrom"])
    dist.barrier(group=group)                                    # nobody overwrites a file rank 0 is still reading
    return out


def bai_reference_spans(bai_path):
    """Per reference of a .bai (SAMv1 5.2): (first virtual offset, last virtual offset) of its records' bin chunks, or None
    for a reference without reads.  What svtk_load_bam_refs uses to find a reference group's bytes in the file."""
    import struct
    b = open(bai_path, "rb").read()
    assert b[:4] == b"BAI\1", "not a BAI index"
    n_ref = struct.unpack_from("<i", b, 4)[0]
    p, out = 8, []
    for _ in range(n_ref):
        n_bin = struct.unpack_from("<i", b, p)[0]
        p += 4
        lo, hi = None, None
        for _ in range(n_bin):
            bin_id, n_chunk = struct.unpack_from("<Ii", b, p)
            p += 8
            if bin_id != 37450:
                for k in range(n_chunk):
                    vb, ve = struct.unpack_from("<QQ", b, p + 16 * k)
                    lo = vb if lo is None else min(lo, vb)
                    hi = ve if hi is None else max(hi, ve)
            p += 16 * n_chunk
        n_intv = struct.unpack_from("<i", b, p)[0]
        p += 4 + 8 * n_intv
        out.append(None if lo is None else (lo, hi))
    return out


def plan_ref_shards(bai_path, n_shards):
    """Contiguous groups of references, one per shard, balanced by the compressed bytes their records occupy in the file
    (north star: records sharded by genomic region, no collective).  Returns [(tid_lo, tid_hi)], possibly fewer than n_shards."""
    spans = bai_reference_spans(bai_path)
    size = [0 if s is None else max(1, (s[1] >> 16) - (s[0] >> 16)) for s in spans]
    total = float(sum(size)) or 1.0
    shards, lo, acc, done = [], 0, 0.0, 0.0
    for t, sz in enumerate(size):
        acc += sz
        left = n_shards - len(shards)
        if left > 1 and acc >= (total - done) / left and t + 1 < len(size):
            shards.append((lo, t)); lo = t + 1; done += acc; acc = 0.0
    if lo < len(size):
        shards.append((lo, len(size) - 1))
    return shards
