"""K3 -- insertion consensus by partial-order alignment -- against abPOA (the reference's submodule, v1.5.4).

SVTrek never calls abPOA, so parity is pinned at the abPOA library boundary: tests/golden/poa_cases.npz
holds inputs and abpoa_msa's outputs (consensus, coverage, and per read best_score / graph size / alignment hash)
written by tools/make_poa_golden.py from the compiled, unmodified library.  Bar: bit-exact, no tolerance.

  not gpu : the golden file agrees with the compiled abPOA where oracle/_ref exists; the kernel source compiled as
            plain C++ with warp width 1 (tests/hostsim) reproduces every golden value -- algorithm check without a GPU
  gpu     : svtk_poa_consensus through the C ABI reproduces every golden value, and fresh random / tandem-repeat /
            ragged / 32-bit-score / scratch-regrowth cases agree with abPOA run live on the box's CPU
"""
import ctypes as C
import os
import subprocess
import tempfile

import numpy as np
import pytest

from svtrek_b200 import simpoa as P

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden", "poa_cases.npz")
KEYS = ("best_score", "node_n", "n_ops", "ops_hash")


def golden():
    z = np.load(GOLD)
    sets = []
    for s in range(len(z["set_off"]) - 1):
        rs = []
        for r in range(z["set_off"][s], z["set_off"][s + 1]):
            rs.append(z["bases"][z["seq_off"][r]:z["seq_off"][r + 1]])
        sets.append(rs)
    return z, sets


@pytest.fixture(scope="module")
def hostsim():
    d = tempfile.mkdtemp(prefix="svtk_hostsim_")
    so = os.path.join(d, "libpoa_hostsim.so")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-o", so, os.path.join(ROOT, "tests", "hostsim", "poa_hostsim.cpp")])
    L = C.CDLL(so)
    V = C.c_void_p
    L.hostsim_poa.argtypes = [C.c_int, V, V, V, C.c_int, V, V, V, V, V, V, C.c_int, V, C.c_int, V, V]
    return L


def run_hostsim(L, seqs):
    n = len(seqs)
    off = np.zeros(n + 1, np.int64); off[1:] = np.cumsum([len(s) for s in seqs])
    b = np.ascontiguousarray(np.concatenate(seqs), dtype=np.uint8)
    cap = int(off[-1]) + 8
    cons = np.zeros(cap, np.uint8); cov = np.zeros(cap, np.int32)
    bs = np.zeros(n, np.int32); nn = np.zeros(n, np.int32); no = np.zeros(n, np.int32); hh = np.zeros(n, np.uint64); cc = np.zeros(n, np.int64)
    st = C.c_int(0); ct = C.c_longlong(0)
    ln = L.hostsim_poa(n, off.ctypes.data, b.ctypes.data, cons.ctypes.data, cap, cov.ctypes.data, bs.ctypes.data, nn.ctypes.data,
                       no.ctypes.data, hh.ctypes.data, cc.ctypes.data, -1, None, 0, C.byref(st), C.byref(ct))
    assert st.value == 0, f"hostsim status {st.value}"
    return dict(cons=cons[:ln], cov=cov[:ln], best_score=bs, node_n=nn, n_ops=no, ops_hash=hh, cells=cc)


def test_golden_file_matches_compiled_abpoa():
    from oracle import binding as B
    if not B.have_abpoa():
        pytest.skip("oracle/_ref/libabpoa_ref.so not built here")
    z, sets = golden()
    for s in (0, 1, 2, 5, 10, 47, 50, 53):
        r = B.ref_poa_steps(sets[s])
        assert np.array_equal(r["cons"], z["cons"][z["cons_off"][s]:z["cons_off"][s + 1]]), z["names"][s]
        lo, hi = z["set_off"][s], z["set_off"][s + 1]
        for k in KEYS:
            assert np.array_equal(r[k], z[k][lo:hi]), (z["names"][s], k)
    # abPOA's own test_data
    assert P.decode(z["cons"][z["cons_off"][0]:z["cons_off"][1]]) == "CGTCAATCTATCGAAGCATACGCGGCAGAGCCGAAGACCTCGGCAATCAC"
    assert P.decode(z["cons"][z["cons_off"][1]:z["cons_off"][2]]) == "ACGTGTACACGTTGAC"


def test_kernel_source_at_warp_width_1_matches_golden(hostsim):
    z, sets = golden()
    for s, seqs in enumerate(sets):
        r = run_hostsim(hostsim, seqs)
        lo, hi = z["set_off"][s], z["set_off"][s + 1]
        assert np.array_equal(r["cons"], z["cons"][z["cons_off"][s]:z["cons_off"][s + 1]]), z["names"][s]
        assert np.array_equal(r["cov"], z["cov"][z["cons_off"][s]:z["cons_off"][s + 1]]), z["names"][s]
        for k in KEYS:
            assert np.array_equal(r[k], z[k][lo:hi]), (z["names"][s], k)
        # band cells: abPOA snaps a row's left edge per SIMD vector (abpoa_align_simd.c:957-959), this path per cell
        assert abs(int(r["cells"].sum()) - int(z["cells"][lo:hi].sum())) <= 0.05 * int(z["cells"][lo:hi].sum()) + 64


def test_pack_layout():
    sets = [[np.array([0, 1, 2], np.uint8)], [np.array([3], np.uint8), np.array([3, 3], np.uint8)]]
    so, qo, b = P.pack(sets)
    assert so.tolist() == [0, 1, 3] and qo.tolist() == [0, 3, 4, 6] and b.tolist() == [0, 1, 2, 3, 3, 3]


# ------------------------------------------------------------------------------------------------------------ GPU
@pytest.mark.gpu
def test_gpu_matches_golden(auditor):
    z, sets = golden()
    cons, st, cov, tr = auditor.poa_consensus(z["set_off"], z["seq_off"], z["bases"], want_cov=True, want_trace=True)
    assert st["n_sets"] == len(sets) and st["cells"] > 0 and st["kernel_ms"] > 0
    for s in range(len(sets)):
        lo, hi = z["set_off"][s], z["set_off"][s + 1]
        assert np.array_equal(cons[s], z["cons"][z["cons_off"][s]:z["cons_off"][s + 1]]), z["names"][s]
        assert np.array_equal(cov[s], z["cov"][z["cons_off"][s]:z["cons_off"][s + 1]]), z["names"][s]
        for k in KEYS:
            assert np.array_equal(tr[k][lo:hi], z[k][lo:hi]), (z["names"][s], k)


@pytest.mark.gpu
def test_gpu_matches_abpoa_live(auditor):
    from oracle import binding as B
    if not B.have_abpoa():
        pytest.skip("oracle/_ref/libabpoa_ref.so did not travel")
    rng = np.random.default_rng(77)
    sets = [P.read_set(rng, ("random", "tandem", "ragged")[t % 3]) for t in range(60)]
    sets += P.ins_workload(5, 40, depth=30, len_hi=2000)
    so, qo, b = P.pack(sets)
    cons, st, tr = auditor.poa_consensus(so, qo, b, want_trace=True)
    for s, seqs in enumerate(sets):
        r = B.ref_poa_steps(seqs)
        assert np.array_equal(cons[s], r["cons"]), s
        for k in KEYS:
            assert np.array_equal(tr[k][so[s]:so[s + 1]], r[k]), (s, k)


@pytest.mark.gpu
def test_gpu_32bit_scores_and_scratch_regrowth(auditor):
    """A graph beyond 16 364 rows switches abPOA to int32 scores (abpoa_align_simd.c:1293).  SVTK_POA_SCALE starts the
    workers from a scratch block 20x too small, so both sets run out of nodes / band cells and are rerun with larger blocks."""
    from oracle import binding as B
    if not B.have_abpoa():
        pytest.skip("oracle/_ref/libabpoa_ref.so did not travel")
    rng = np.random.default_rng(5)
    tpl = rng.integers(0, 4, 7000).astype(np.uint8)
    big = [P.mutate(rng, tpl, 0.15) for _ in range(20)]           # 7 kb at 15 %: > 16 364 nodes from the 16th read on
    unrelated = [rng.integers(0, 4, 300).astype(np.uint8) for _ in range(30)]
    sets = [big, unrelated]
    so, qo, b = P.pack(sets)
    os.environ["SVTK_POA_SCALE"] = "0.05"
    try:
        cons, st, tr = auditor.poa_consensus(so, qo, b, want_trace=True)
    finally:
        del os.environ["SVTK_POA_SCALE"]
    assert st["relaunched"] >= 1
    for s, seqs in enumerate(sets):
        r = B.ref_poa_steps(seqs)
        assert np.array_equal(cons[s], r["cons"]), s
        for k in KEYS:
            assert np.array_equal(tr[k][so[s]:so[s + 1]], r[k]), (s, k)
    assert tr["node_n"][so[1] - 2] > 16364                      # the last alignment of `big` ran on int32 scores


@pytest.mark.gpu
def test_gpu_rejects_empty_reads_and_handles_empty_batch(auditor):
    from svtrek_b200 import api
    cons, st = auditor.poa_consensus([0], [0], np.zeros(0, np.uint8))
    assert cons == [] and st["n_sets"] == 0
    with pytest.raises(api.SvtkError):
        auditor.poa_consensus([0, 2], [0, 3, 3], np.array([0, 1, 2], np.uint8))
    # idempotence / order independence of the batch: the same set twice and in a different batch position
    rng = np.random.default_rng(3)
    a, bset = P.read_set(rng, "random", 300, 12, 0.1), P.read_set(rng, "tandem")
    so, qo, b = P.pack([a, bset, a])
    cons, _ = auditor.poa_consensus(so, qo, b)
    assert np.array_equal(cons[0], cons[2])
    so2, qo2, b2 = P.pack([bset, a])
    cons2, _ = auditor.poa_consensus(so2, qo2, b2)
    assert np.array_equal(cons2[0], cons[1]) and np.array_equal(cons2[1], cons[0])
