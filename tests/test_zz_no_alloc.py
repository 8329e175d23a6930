"""svtk_consensus_pos_batch_static (include/svtk.h): consensus_pos on device arrays with the sort scratch in a file-scope device
array -- the entry point for callers that cannot call cudaMalloc.  It has no counterpart in the reference; its results must be
those of svtk_consensus_pos_batch (refinement.c:41 through the ctx arena) and of the CPU oracle, bit for bit, on two calls in a
row (the second one finds the first one's contents in the shared array).  Named test_zz_* so that pytest collects it last."""
import numpy as np
import pytest

from oracle import binding as B
from svtrek_b200 import api

pytestmark = pytest.mark.gpu


def make_case(seed, n_seg, long_every, pos_hi=2_000_000):
    """Segments of mixed length: empty, below min-count, shared-memory sized, and (every `long_every`-th) longer than the 8 192
    ints that fit the CTA's shared segment, i.e. sorted in the static scratch.  Values cluster around pos like signatures do."""
    rng = np.random.default_rng(seed)
    lens = rng.integers(0, 60, n_seg)
    lens[::7] = rng.integers(0, 3, len(lens[::7]))
    lens[3::long_every] = rng.integers(8193, 40000, len(lens[3::long_every]))
    off = np.zeros(n_seg + 1, np.int64)
    off[1:] = np.cumsum(lens)
    pos = rng.integers(10_000, pos_hi, n_seg).astype(np.int32)
    vals = np.empty(int(off[-1]), np.int32)
    for s in range(n_seg):
        n = int(lens[s])
        near = rng.random(n) < 0.6
        v = np.where(near, pos[s] + rng.integers(-700, 700, n), rng.integers(0, pos_hi, n))
        if s % 3:                                           # a tight cluster near pos, as real signatures form: the vote has work to do
            k = n // 3 if n > 8192 else min(n, int(rng.integers(2, 9)))
            v[:k] = pos[s] + rng.integers(-400, 400) + rng.integers(-3, 4, k)
        vals[off[s]:off[s + 1]] = v
    return vals, off, pos


def run_static(a, torch, vals, off, pos):
    dev = torch.device("cuda:0")
    d_val = torch.from_numpy(vals).to(dev)
    d_off = torch.from_numpy(off).to(dev)
    d_pos = torch.from_numpy(pos).to(dev)
    d_res = torch.full((len(pos),), 12345, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()                                # the ctx has its own stream
    a.consensus_pos_batch_static(d_val.data_ptr(), d_off.data_ptr(), len(pos), len(vals), d_pos.data_ptr(), d_res.data_ptr())
    a.sync()
    return d_res.cpu().numpy()


def test_static_scratch_variant_equals_the_allocating_entry_point_twice_in_a_row():
    torch = pytest.importorskip("torch")
    p = api.Params(**api.DEFAULT_PARAMS)
    with api.Auditor(0) as a:
        first = None
        for call, (seed, n_seg, long_every) in enumerate([(11, 600, 40), (12, 900, 30)]):
            vals, off, pos = make_case(seed, n_seg, long_every)
            assert int((np.diff(off) > 8192).sum()) >= 10 and len(vals) <= api.STATIC_MAX_VALUES
            want = a.consensus_pos_batch(vals, off, pos)                    # existing entry point (ctx arena)
            got = run_static(a, torch, vals, off, pos)
            assert np.array_equal(got, want), f"call {call}: {int((got != want).sum())} segments differ"    # integer work: bit-exact
            orc = np.array([B.consensus_pos(vals[off[s]:off[s + 1]], int(pos[s]), p.consensus_min_count, p.consensus_interval,
                                            p.consensus_interval_range) if off[s + 1] - off[s] >= p.consensus_min_count else -1
                            for s in range(0, n_seg, 9)], np.int32)
            assert np.array_equal(got[::9], orc)
            assert (got >= 0).sum() > n_seg // 5                            # not a table of NAs
            if first is None:
                first = (vals, off, pos, got)
        # the first case again, after the second call has overwritten the shared array
        vals, off, pos, got0 = first
        assert np.array_equal(run_static(a, torch, vals, off, pos), got0)


def test_static_scratch_variant_rejects_what_does_not_fit():
    torch = pytest.importorskip("torch")
    with api.Auditor(0) as a:
        d = torch.zeros(8, dtype=torch.int64, device="cuda:0")
        with pytest.raises(api.SvtkError, match="SVTK_STATIC_MAX_VALUES"):
            a.consensus_pos_batch_static(d.data_ptr(), d.data_ptr(), 1, api.STATIC_MAX_VALUES + 1, d.data_ptr(), d.data_ptr())
        # a segment that leaves [0, n_values] is flagged on the device and touches nothing
        vals = np.arange(100, dtype=np.int32)
        off = np.array([0, 50, 200], np.int64)
        pos = np.array([25, 75], np.int32)
        got = run_static(a, torch, vals, off, pos)
        assert got[1] == api.STATIC_BAD_SEGMENT and got[0] == a.consensus_pos(vals[:50], 25)
