"""Region sharding: shards are closed under their reads, results merge back in VCF order, and the
world-size-2 path runs over torch.distributed (gloo on CPU; the refine step of each rank is played by the CPU oracle here,
the CUDA path takes its place on the GPU box)."""
import os
import socket

import numpy as np
import pytest

from svtrek_b200 import shard, simreads as S


def _oracle_refine(sub, chrom, start, end, pos, kind):
    from oracle import binding as B
    return B.refine_batch(B.OracleReads(sub), chrom, start, end, pos, kind)


def _job(seed=11):
    rs, truth = S.simulate(seed, [400_000, 300_000], 20, 12000, 0.005, 60, margin=30000)
    return rs, S.windows_from_truth(truth)


def test_plan_keeps_records_together_and_covers_everything():
    rs, w = _job()
    for n in (1, 2, 3, 8):
        sh = shard.plan_shards(w, n)
        assert sh[0][0] == 0 and sh[-1][1] == len(w["chrom"])
        for (a, b), (c, d) in zip(sh[:-1], sh[1:]):
            assert b == c
            if 0 < b < len(w["chrom"]):
                assert w["record"][b] != w["record"][b - 1]


@pytest.mark.parametrize("n", [2, 3, 5])
def test_shards_are_closed_under_their_reads(n):
    rs, w = _job()
    whole = _oracle_refine(rs, w["chrom"], w["start"], w["end"], w["pos"], w["kind"])
    parts = []
    for lo, hi in shard.plan_shards(w, n):
        sub = shard.shard_reads(rs, w, lo, hi)
        assert sub.n_reads <= rs.n_reads
        parts.append(_oracle_refine(sub, *[w[k][lo:hi] for k in ("chrom", "start", "end", "pos", "kind")]))
    assert np.array_equal(np.concatenate(parts), whole)
    assert (whole != -1).sum() > 10


def _worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rs, w = _job()
    res = shard.audit_sharded(_oracle_refine, rs, w, rank, world)
    q.put((rank, res))
    dist.barrier()
    dist.destroy_process_group()


def test_world_size_2_gloo_merge_in_vcf_order():
    import torch.multiprocessing as mp
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=300) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    rs, w = _job()
    whole = _oracle_refine(rs, w["chrom"], w["start"], w["end"], w["pos"], w["kind"])
    assert np.array_equal(got[0], whole) and got[1] is None          # merged on rank 0's host, from the shard files


@pytest.mark.gpu
def test_gpu_sharded_equals_whole(auditor):
    rs, w = _job(13)
    auditor.load_reads(rs)
    whole = auditor.refine_windows(w["chrom"], w["start"], w["end"], w["pos"], w["kind"])

    def refine(sub, chrom, start, end, pos, kind):
        auditor.load_reads(sub)
        return auditor.refine_windows(chrom, start, end, pos, kind)
    parts = [shard.audit_sharded(refine, shard.shard_reads(rs, w, lo, hi), {k: v[lo:hi] for k, v in w.items()}, 0, 1)
             for lo, hi in shard.plan_shards(w, 4, rs)]
    assert np.array_equal(np.concatenate(parts), whole)


def test_plan_ref_shards_on_the_golden_index():
    import os
    from svtrek_b200 import shard as SH
    bai = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "bam_small.bam.bai")
    spans = SH.bai_reference_spans(bai)
    assert len(spans) == 3 and spans[1] is None and spans[0][0] < spans[0][1] <= spans[2][0] < spans[2][1]
    for n in (1, 2, 3, 5):
        sh = SH.plan_ref_shards(bai, n)
        assert 1 <= len(sh) <= n and sh[0][0] == 0 and sh[-1][1] == 2
        assert all(a[1] + 1 == b[0] for a, b in zip(sh, sh[1:]))
