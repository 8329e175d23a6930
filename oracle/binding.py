"""TEST INFRASTRUCTURE ONLY: ctypes bindings to (a) the C oracle (oracle/svtk_oracle.c) and (b) the compiled,
unmodified reference (oracle/_ref/libsvtrek_ref.so + oracle/_ref/svtrek, built by oracle/build_ref.sh).

Importers: tests/, __graft_entry__.smoke(), bench.py's cpu_baseline and --impl reference legs. Nothing under
svtrek_b200/ may import this module."""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")
KIND_DEL_START, KIND_DEL_END, KIND_INS, KIND_POINT = 0, 1, 2, 3


class _Reads(C.Structure):
    _fields_ = [("n_tid", C.c_int32), ("tid_off", C.c_void_p), ("pos", C.c_void_p), ("end", C.c_void_p),
                ("n_cigar", C.c_void_p), ("cigar_off", C.c_void_p), ("cigar", C.c_void_p)]


class _Params(C.Structure):
    _fields_ = [("wider", C.c_int), ("median", C.c_int), ("narrow", C.c_int), ("range", C.c_int),
                ("interval", C.c_int), ("min_count", C.c_int)]


DEFAULT_PARAMS = dict(wider=20000, median=10000, narrow=2000, range=500, interval=5, min_count=3)


def build_oracle():
    subprocess.check_call(["make", "-s", "-C", HERE, os.path.join(HERE, "libsvtk_oracle.so")])


_lib = None


def lib():
    global _lib
    if _lib is None:
        path = os.path.join(HERE, "libsvtk_oracle.so")
        if not os.path.exists(path) or os.path.getmtime(path) < os.path.getmtime(os.path.join(HERE, "svtk_oracle.c")):
            build_oracle()
        L = C.CDLL(path)
        L.orc_consensus_pos.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]
        L.orc_consensus_pos.restype = C.c_int
        L.orc_lower_bound.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.orc_upper_bound.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.orc_refine.argtypes = [C.POINTER(_Reads), C.c_int, C.c_int, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int, C.c_int,
                                 C.c_int, C.c_void_p, C.c_int, C.POINTER(C.c_int)]
        L.orc_refine.restype = C.c_int
        L.orc_sliding_window_ins.argtypes = [C.POINTER(_Reads), C.c_int, C.c_uint32, C.c_uint32, C.c_int, C.c_int, C.c_int, C.c_void_p]
        L.orc_sliding_window_ins.restype = C.c_int
        L.orc_sliding_window_ins_path.argtypes = [C.POINTER(_Reads), C.c_int, C.c_uint32, C.c_uint32, C.c_int, C.c_int, C.c_int, C.c_char_p]
        L.orc_sliding_window_ins_path.restype = C.c_int
        L.orc_audit_file.argtypes = [C.POINTER(_Reads), C.c_char_p, C.POINTER(_Params), C.c_char_p, C.c_char_p]
        L.orc_audit_file.restype = C.c_long
        L.orc_refine_batch.argtypes = [C.POINTER(_Reads), C.c_int] + [C.c_void_p] * 5 + [C.c_int] * 3 + [C.c_void_p]
        L.orc_refine_batch.restype = None
        _lib = L
    return _lib


class OracleReads:
    """Keeps the numpy arrays alive and exposes them as the oracle's SoA struct."""

    def __init__(self, rs):
        self.arrs = [np.ascontiguousarray(rs.tid_off, dtype=np.int64), np.ascontiguousarray(rs.pos, dtype=np.uint32),
                     np.ascontiguousarray(rs.end, dtype=np.uint32), np.ascontiguousarray(rs.n_cigar, dtype=np.uint32),
                     np.ascontiguousarray(rs.cigar_off, dtype=np.uint64), np.ascontiguousarray(rs.cigar, dtype=np.uint32)]
        self.c = _Reads(int(self.arrs[0].shape[0] - 1), *[a.ctypes.data for a in self.arrs])


def consensus_pos(values, pos, min_count=3, interval=5, range_=500):
    a = np.array(values, dtype=np.int32)
    return lib().orc_consensus_pos(a.ctypes.data, int(a.shape[0]), int(pos), int(min_count), int(interval), int(range_))


def lower_bound(values, x):
    a = np.ascontiguousarray(values, dtype=np.int32)
    return lib().orc_lower_bound(a.ctypes.data, int(a.shape[0]), int(x))


def upper_bound(values, x):
    a = np.ascontiguousarray(values, dtype=np.int32)
    return lib().orc_upper_bound(a.ctypes.data, int(a.shape[0]), int(x))


def refine(oreads, kind, chrom, start, end, pos, min_count=3, interval=5, range_=500, want_sig=False, cap=1 << 16):
    n = C.c_int(0)
    sig = np.zeros(cap, dtype=np.int32) if want_sig else None
    r = lib().orc_refine(C.byref(oreads.c), kind, chrom, start & 0xFFFFFFFF, end & 0xFFFFFFFF, pos & 0xFFFFFFFF, min_count,
                         interval, range_, sig.ctypes.data if want_sig else None, cap, C.byref(n))
    if want_sig:
        return r, np.sort(sig[:min(n.value, cap)])
    return r


def refine_batch(oreads, chrom, w_start, w_end, pos, kind, min_count=3, interval=5, range_=500):
    n = len(chrom)
    chrom = np.ascontiguousarray(chrom, dtype=np.int32); w_start = np.ascontiguousarray(w_start, dtype=np.uint32)
    w_end = np.ascontiguousarray(w_end, dtype=np.uint32); pos = np.ascontiguousarray(pos, dtype=np.uint32)
    kind = np.ascontiguousarray(kind, dtype=np.uint8)
    out = np.zeros(n, dtype=np.int32)
    lib().orc_refine_batch(C.byref(oreads.c), n, chrom.ctypes.data, w_start.ctypes.data, w_end.ctypes.data, pos.ctypes.data,
                           kind.ctypes.data, min_count, interval, range_, out.ctypes.data)
    return out


def sliding_window_ins(oreads, chrom, start, end, min_count, window, slide, want_lines=False):
    if not want_lines:
        return lib().orc_sliding_window_ins(C.byref(oreads.c), chrom, start, end, min_count, window, slide, None)
    with tempfile.TemporaryDirectory() as d:
        o = os.path.join(d, "o")
        r = lib().orc_sliding_window_ins_path(C.byref(oreads.c), chrom, start, end, min_count, window, slide, o.encode())
        return r, (open(o).read() if os.path.exists(o) else "")


def audit_file(oreads, vcf_path, params=None, want_err=False):
    """Oracle restatement of `svtrek audt -t 1` stdout (and stderr) for a VCF file over an SoA read set."""
    p = dict(DEFAULT_PARAMS); p.update(params or {})
    cp = _Params(p["wider"], p["median"], p["narrow"], p["range"], p["interval"], p["min_count"])
    with tempfile.TemporaryDirectory() as d:
        o, e = os.path.join(d, "o"), os.path.join(d, "e")
        n = lib().orc_audit_file(C.byref(oreads.c), vcf_path.encode(), C.byref(cp), o.encode(), e.encode())
        if n < 0:
            raise RuntimeError("oracle audit failed")
        out = open(o).read()
        err = open(e).read()
    return (out, err) if want_err else out


# ------------------------------------------------------------------------------------------------------------
# The compiled reference (only where oracle/_ref exists: this container after build_ref.sh, and the GPU box
# because the built files travel with the snapshot).
def have_ref():
    return os.path.exists(os.path.join(REF_DIR, "libsvtrek_ref.so")) and os.path.exists(os.path.join(REF_DIR, "svtrek"))


_ref = None


def ref():
    global _ref
    if _ref is None:
        L = C.CDLL(os.path.join(REF_DIR, "libsvtrek_ref.so"))
        L.ref_consensus_pos.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]
        L.ref_lower_bound.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.ref_upper_bound.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.ref_open.argtypes = [C.c_char_p]; L.ref_open.restype = C.c_void_p
        L.ref_close.argtypes = [C.c_void_p]
        L.ref_refine.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int, C.c_int, C.c_int,
                                 C.c_void_p, C.c_int, C.POINTER(C.c_int)]
        L.ref_sliding_window_ins.argtypes = [C.c_void_p, C.c_int, C.c_uint32, C.c_uint32, C.c_int, C.c_int, C.c_int]
        _ref = L
    return _ref


def ref_consensus_pos(values, pos, min_count=3, interval=5, range_=500):
    a = np.array(values, dtype=np.int32)
    return ref().ref_consensus_pos(a.ctypes.data, int(a.shape[0]), int(pos), int(min_count), int(interval), int(range_))


class RefBam:
    def __init__(self, bam_path):
        self.h = ref().ref_open(bam_path.encode())
        if not self.h:
            raise RuntimeError(f"reference could not open {bam_path}")

    def refine(self, kind, chrom, start, end, pos, min_count=3, interval=5, range_=500, want_sig=False, cap=1 << 16):
        n = C.c_int(0)
        sig = np.zeros(cap, dtype=np.int32) if want_sig else None
        r = ref().ref_refine(self.h, kind, chrom, start & 0xFFFFFFFF, end & 0xFFFFFFFF, pos & 0xFFFFFFFF, min_count, interval,
                             range_, sig.ctypes.data if want_sig else None, cap, C.byref(n) if want_sig else None)
        if want_sig:
            return r, np.sort(sig[:min(n.value, cap)])
        return r

    def sliding_window_ins(self, chrom, start, end, min_count, window, slide, want_lines=False):
        """The reference prints its per-tile lines with printf: capture fd 1 around the call."""
        if not want_lines:
            return ref().ref_sliding_window_ins(self.h, chrom, start, end, min_count, window, slide)
        import sys
        sys.stdout.flush()
        with tempfile.TemporaryFile() as tf:
            saved = os.dup(1)
            os.dup2(tf.fileno(), 1)
            try:
                r = ref().ref_sliding_window_ins(self.h, chrom, start, end, min_count, window, slide)
            finally:
                os.dup2(saved, 1)
                os.close(saved)
            tf.seek(0)
            return r, tf.read().decode()

    def close(self):
        if self.h:
            ref().ref_close(self.h)
            self.h = None


def make_bam(sam_path, bam_path):
    """SAM -> BAM + BAI with HTSlib's own helper tools (no samtools on the image)."""
    subprocess.check_call([os.path.join(REF_DIR, "test_view"), "-b", "-p", bam_path, sam_path], stdout=subprocess.DEVNULL)
    subprocess.check_call([os.path.join(REF_DIR, "test_index"), "-b", bam_path], stdout=subprocess.DEVNULL)


def run_ref_cli(bam, vcf, threads=1, extra=(), capture=True, timeout=3600):
    """`./svtrek audt -b BAM -v VCF -t N` -- the reference binary, unmodified. Returns (stdout, stderr)."""
    cmd = [os.path.join(REF_DIR, "svtrek"), "audt", "-b", bam, "-v", vcf, "-t", str(threads), *map(str, extra)]
    if capture:
        p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, timeout=timeout, check=True)
        return p.stdout.decode(), p.stderr.decode()
    subprocess.run(cmd, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, timeout=timeout, check=True)
    return None, None


# ------------------------------------------------------------------------------------------------------------
# K3 oracle: the compiled, unmodified abPOA library behind oracle/poa_harness.c (oracle/_ref/libabpoa_ref.so).
def have_abpoa():
    return os.path.exists(os.path.join(REF_DIR, "libabpoa_ref.so"))


_abpoa = None


def abpoa():
    global _abpoa
    if _abpoa is None:
        L = C.CDLL(os.path.join(REF_DIR, "libabpoa_ref.so"))
        V = C.c_void_p
        L.ref_poa_msa.argtypes = [C.c_int, V, V, V, C.c_int, V]
        L.ref_poa_steps.argtypes = [C.c_int, V, V, V, V, V, V, V, V, C.c_int, V, C.c_int, V, C.c_int]
        _abpoa = L
    return _abpoa


def ref_poa_msa(seqs):
    """abpoa_msa (defaults) on one read set: list of uint8 code arrays -> (consensus codes, coverage)."""
    lens = [len(s) for s in seqs]
    off = np.zeros(len(seqs) + 1, np.int64); off[1:] = np.cumsum(lens)
    b = np.ascontiguousarray(np.concatenate(seqs), dtype=np.uint8)
    cap = int(off[-1]) + 8
    cons = np.zeros(cap, np.uint8); cov = np.zeros(cap, np.int32)
    n = abpoa().ref_poa_msa(len(seqs), off.ctypes.data, b.ctypes.data, cons.ctypes.data, cap, cov.ctypes.data)
    return cons[:n].copy(), cov[:n].copy()


def ref_poa_steps(seqs):
    """The same default path read by read: consensus, coverage and per-read best_score / node_n / n_ops / ops_hash / cells."""
    n = len(seqs)
    lens = [len(s) for s in seqs]
    off = np.zeros(n + 1, np.int64); off[1:] = np.cumsum(lens)
    b = np.ascontiguousarray(np.concatenate(seqs), dtype=np.uint8)
    cap = int(off[-1]) + 8
    cons = np.zeros(cap, np.uint8); cov = np.zeros(cap, np.int32)
    bs = np.zeros(n, np.int32); nn = np.zeros(n, np.int32); no = np.zeros(n, np.int32); hh = np.zeros(n, np.uint64); cc = np.zeros(n, np.int64)
    ln = abpoa().ref_poa_steps(n, off.ctypes.data, b.ctypes.data, bs.ctypes.data, nn.ctypes.data, no.ctypes.data, hh.ctypes.data,
                               cc.ctypes.data, cons.ctypes.data, cap, cov.ctypes.data, -1, None, 0)
    return dict(cons=cons[:ln].copy(), cov=cov[:ln].copy(), best_score=bs, node_n=nn, n_ops=no, ops_hash=hh, cells=cc)


# ---- inserted sequences through the reference's HTSlib (oracle/ins_harness.c -> oracle/_ref/libins_ref.so)
def have_ins_ref():
    return os.path.exists(os.path.join(REF_DIR, "libins_ref.so"))


def ref_ins_sequences(bam_path, chrom, start, end, kind, refined, tol=5, min_len=50):
    """Per INS window with a refined position: the bases of every `I >= min_len` op refine_ins' loop meets within tol of it, read
    with sam_itr_queryi / sam_itr_next / bam_seqi of the reference's HTSlib.  Returns dict(window, read_pos, ref_pos, seq_off, bases)."""
    L = C.CDLL(os.path.join(REF_DIR, "libins_ref.so"))
    L.ref_ins_sequences.argtypes = [C.c_char_p, C.c_int] + [C.c_void_p] * 5 + [C.c_int, C.c_int, C.c_int64, C.c_int64] + [C.c_void_p] * 5 + [C.POINTER(C.c_int64)] * 2
    chrom = np.ascontiguousarray(chrom, dtype=np.int32)
    start = np.ascontiguousarray(np.asarray(start).astype(np.int64) & 0xFFFFFFFF, dtype=np.uint32)
    end = np.ascontiguousarray(np.asarray(end).astype(np.int64) & 0xFFFFFFFF, dtype=np.uint32)
    kind = np.ascontiguousarray(kind, dtype=np.uint8)
    refined = np.ascontiguousarray(refined, dtype=np.int32)
    cap_s, cap_b = 1 << 12, 1 << 20
    while True:
        win = np.zeros(cap_s, np.int32); rp = np.zeros(cap_s, np.uint32); xp = np.zeros(cap_s, np.uint32)
        off = np.zeros(cap_s + 1, np.int64); bases = np.zeros(cap_b, np.uint8)
        ns, nb = C.c_int64(0), C.c_int64(0)
        rc = L.ref_ins_sequences(os.fsencode(bam_path), int(len(chrom)), chrom.ctypes.data, start.ctypes.data, end.ctypes.data, kind.ctypes.data,
                                 refined.ctypes.data, int(tol), int(min_len), cap_s, cap_b, win.ctypes.data, rp.ctypes.data, xp.ctypes.data,
                                 off.ctypes.data, bases.ctypes.data, C.byref(ns), C.byref(nb))
        if rc == -2:
            raise IOError("ref_ins_sequences: cannot open " + str(bam_path))
        if rc == 0:
            n = int(ns.value)
            return dict(window=win[:n].copy(), read_pos=rp[:n].copy(), ref_pos=xp[:n].copy(), seq_off=off[:n + 1].copy(), bases=bases[:int(nb.value)].copy())
        cap_s, cap_b = max(cap_s, int(ns.value) + 1), max(cap_b, int(nb.value) + 1)
