"""ctypes mirror of include/svtk.h -- the host-side face of the B200 audt refinement path.

Names follow the reference's own (refinement.h / audit.c): deletion, insertion, inversion, consensus_pos,
sliding_window_ins, process_vcf.  Everything here calls into libsvtk_b200.so (hand-written sm_100a kernels);
there is no Python or CPU fallback -- a missing library or a missing GPU raises."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("SVTK_LIB") or os.path.join(HERE, "libsvtk_b200.so")    # SVTK_LIB: kernel experiments only

DEL_START, DEL_END, INS, POINT = 0, 1, 2, 3
DEFAULT_PARAMS = dict(wider_interval=20000, median_interval=10000, narrow_interval=2000,
                      consensus_interval_range=500, consensus_interval=5, consensus_min_count=3)   # params.h:27-32


class SvtkError(RuntimeError):
    pass


class Interval(C.Structure):
    _fields_ = [("start", C.c_uint32), ("end", C.c_uint32)]


class Params(C.Structure):
    _fields_ = [(k, C.c_int32) for k in DEFAULT_PARAMS]


class Reads(C.Structure):
    _fields_ = [("n_tid", C.c_int32), ("n_reads", C.c_int64), ("n_words", C.c_int64), ("tid_off", C.c_void_p),
                ("pos", C.c_void_p), ("end", C.c_void_p), ("n_cigar", C.c_void_p), ("cigar_off", C.c_void_p), ("cigar", C.c_void_p)]


class Windows(C.Structure):
    _fields_ = [("n_win", C.c_int32), ("chrom", C.c_void_p), ("start", C.c_void_p), ("end", C.c_void_p), ("pos", C.c_void_p),
                ("kind", C.c_void_p)]


class Stats(C.Structure):
    _fields_ = [(k, C.c_uint64) for k in ("n_windows", "n_pairs", "n_words_visited", "n_signatures", "n_overflow", "algorithmic_bytes")]


class Timing(C.Structure):
    _fields_ = [("refine_kernel_ms", C.c_double), ("refine_kernel_launches", C.c_uint64), ("index_kernel_ms", C.c_double),
                ("index_kernel_launches", C.c_uint64), ("h2d_bytes", C.c_uint64), ("d2h_bytes", C.c_uint64)]


class IndexStats(C.Structure):
    _fields_ = [(k, C.c_uint64) for k in ("n_reads", "n_words", "n_chunks", "n_events", "n_segments", "index_bytes", "n_windows",
                                          "refine_events_scanned", "refine_segments", "refine_reads_in_range", "n_signatures",
                                          "n_overflow", "refine_bytes")]


class PoaBatch(C.Structure):
    _fields_ = [("n_sets", C.c_int32), ("set_off", C.c_void_p), ("seq_off", C.c_void_p), ("bases", C.c_void_p)]


class PoaTrace(C.Structure):
    _fields_ = [("best_score", C.c_void_p), ("node_n", C.c_void_p), ("n_ops", C.c_void_p), ("ops_hash", C.c_void_p), ("cells", C.c_void_p)]


class PoaStats(C.Structure):
    _fields_ = [(k, C.c_uint64) for k in ("n_sets", "n_reads", "n_bases", "cells", "cells_full", "relaunched", "workers")] + [("kernel_ms", C.c_double), ("phase_cycles", C.c_uint64 * 4)]


class BamStats(C.Structure):
    _fields_ = [("file_bytes", C.c_uint64), ("n_blocks", C.c_uint64), ("inflated_bytes", C.c_uint64), ("n_reads", C.c_uint64), ("n_words", C.c_uint64), ("image_bytes", C.c_uint64),
                ("n_ref", C.c_int32), ("n_pieces", C.c_uint32), ("lanes_per_warp", C.c_uint32), ("host_scan_ms", C.c_float), ("upload_ms", C.c_float), ("inflate_ms", C.c_float),
                ("parse_ms", C.c_float), ("total_ms", C.c_float), ("load_span_ms", C.c_float), ("n_spans", C.c_uint32), ("inflate_launches", C.c_uint32)]


class InsSets(C.Structure):
    _fields_ = [("n_sets", C.c_int32), ("n_seqs", C.c_int64), ("set_off", C.POINTER(C.c_int64)), ("seq_off", C.POINTER(C.c_int64)),
                ("bases", C.POINTER(C.c_uint8)), ("window", C.POINTER(C.c_int32)), ("read", C.POINTER(C.c_int32)), ("ref_pos", C.POINTER(C.c_uint32))]


EXPORTS = ["svtk_version", "svtk_last_error", "svtk_device_count", "svtk_create", "svtk_destroy", "svtk_set_params", "svtk_get_params",
           "svtk_load_reads", "svtk_attach_reads_device", "svtk_refine_windows", "svtk_refine_windows_device", "svtk_measure_windows",
           "svtk_sync", "svtk_launch_count", "svtk_set_profiling", "svtk_get_timing", "svtk_index_reads", "svtk_measure_index", "svtk_deletion", "svtk_insertion", "svtk_inversion", "svtk_consensus_pos_batch",
           "svtk_window_signatures", "svtk_sliding_window_ins", "svtk_process_vcf", "svtk_poa_consensus",
           "svtk_load_bam", "svtk_bam_n_ref", "svtk_bam_ref", "svtk_copy_reads", "svtk_ins_sequences", "svtk_process_vcf_ins", "svtk_load_bam_refs", "svtk_audit_device", "svtk_process_vcf_region", "svtk_vcf_windows", "svtk_vcf_format", "svtk_load_bam_windows", "svtk_process_vcf_flags", "svtk_audit_shard", "svtk_consensus_pos_batch_static"]

STATIC_MAX_VALUES = 1 << 20            # SVTK_STATIC_MAX_VALUES (include/svtk.h)
STATIC_BAD_SEGMENT = -(1 << 31)        # SVTK_STATIC_BAD_SEGMENT

_lib = None


def load_library():
    """dlopen libsvtk_b200.so (built in-tree by svtrek_b200.build). Raises if it is missing: no fallback."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise SvtkError(f"{LIB_PATH} is missing -- run `python -m svtrek_b200.build` (there is no CPU fallback)")
    L = C.CDLL(LIB_PATH)
    L.svtk_last_error.restype = C.c_char_p
    L.svtk_create.restype = C.c_void_p
    L.svtk_create.argtypes = [C.c_int, C.POINTER(Params), C.c_void_p]
    L.svtk_destroy.argtypes = [C.c_void_p]
    L.svtk_set_params.argtypes = [C.c_void_p, C.POINTER(Params)]
    L.svtk_get_params.argtypes = [C.c_void_p, C.POINTER(Params)]
    L.svtk_load_reads.argtypes = [C.c_void_p, C.POINTER(Reads)]
    L.svtk_attach_reads_device.argtypes = [C.c_void_p, C.POINTER(Reads)]
    L.svtk_refine_windows.argtypes = [C.c_void_p, C.POINTER(Windows), C.c_void_p]
    L.svtk_refine_windows_device.argtypes = [C.c_void_p, C.POINTER(Windows), C.c_void_p]
    L.svtk_measure_windows.argtypes = [C.c_void_p, C.POINTER(Windows), C.c_void_p, C.POINTER(Stats)]
    L.svtk_sync.argtypes = [C.c_void_p]
    L.svtk_launch_count.argtypes = [C.c_void_p]
    L.svtk_launch_count.restype = C.c_uint64
    L.svtk_set_profiling.argtypes = [C.c_void_p, C.c_int]
    L.svtk_get_timing.argtypes = [C.c_void_p, C.POINTER(Timing)]
    L.svtk_index_reads.argtypes = [C.c_void_p]
    L.svtk_measure_index.argtypes = [C.c_void_p, C.POINTER(Windows), C.c_void_p, C.POINTER(IndexStats)]
    L.svtk_deletion.argtypes = [C.c_void_p, C.c_int, Interval, Interval, Interval, C.POINTER(Interval)]
    L.svtk_insertion.argtypes = [C.c_void_p, C.c_int, Interval, C.c_uint32, C.POINTER(C.c_uint32)]
    L.svtk_inversion.argtypes = [C.c_void_p, C.c_int, Interval, Interval, Interval, C.POINTER(Interval)]
    L.svtk_consensus_pos_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p]
    L.svtk_consensus_pos_batch_static.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int64, C.c_void_p, C.c_void_p]
    L.svtk_window_signatures.argtypes = [C.c_void_p, C.c_int, C.c_int, Interval, C.c_void_p, C.c_int64, C.POINTER(C.c_int64)]
    L.svtk_sliding_window_ins.argtypes = [C.c_void_p, C.c_int, Interval, C.c_int, C.c_int, C.POINTER(C.c_int32), C.c_char_p, C.c_size_t]
    L.svtk_process_vcf.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.POINTER(C.c_int64)]
    L.svtk_poa_consensus.argtypes = [C.c_void_p, C.POINTER(PoaBatch), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                     C.POINTER(PoaTrace), C.POINTER(PoaStats)]
    L.svtk_load_bam.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.POINTER(BamStats)]
    L.svtk_load_bam_refs.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_int32, C.c_int32, C.POINTER(BamStats)]
    L.svtk_load_bam_windows.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.POINTER(Windows), C.POINTER(BamStats)]
    L.svtk_audit_shard.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_char_p, C.c_char_p, C.c_int32, C.c_int32, C.c_uint32, C.POINTER(C.c_int64), C.POINTER(BamStats)]
    L.svtk_process_vcf_flags.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_char_p, C.c_int32, C.c_uint32, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
    L.svtk_bam_n_ref.argtypes = [C.c_void_p]
    L.svtk_bam_ref.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_char_p), C.POINTER(C.c_uint32)]
    L.svtk_process_vcf_ins.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_char_p, C.c_int32, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
    L.svtk_ins_sequences.argtypes = [C.c_void_p, C.POINTER(Windows), C.c_void_p, C.c_int32, C.c_int32, C.POINTER(InsSets)]
    L.svtk_audit_device.argtypes = [C.c_void_p, C.POINTER(Windows), C.c_void_p]
    L.svtk_process_vcf_region.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_int32, C.c_int32, C.c_uint32, C.POINTER(C.c_int64)]
    L.svtk_vcf_windows.argtypes = [C.c_char_p, C.POINTER(Params), C.c_int32, C.c_int32, C.c_int32, C.c_uint32, C.c_int64] + [C.c_void_p] * 5 + [C.POINTER(C.c_int64)]
    L.svtk_vcf_format.argtypes = [C.c_char_p, C.POINTER(Params), C.c_int32, C.c_int32, C.c_int32, C.c_uint32, C.c_void_p, C.c_int64, C.c_char_p, C.POINTER(C.c_int64)]
    L.svtk_copy_reads.argtypes = [C.c_void_p, C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_int64)] + [C.c_void_p] * 6
    _lib = L
    return L


def _check(rc):
    if rc != 0:
        raise SvtkError(f"svtk error {rc}: {load_library().svtk_last_error().decode()}")


def _u32(a):
    return np.ascontiguousarray(np.asarray(a).astype(np.int64) & 0xFFFFFFFF, dtype=np.uint32)


def vcf_windows(vcf_path, params=None, n_ref=0, refs=None, owns_unplaced=True):
    """svtk_vcf_windows: the refine_* windows process_vcf builds for a VCF (no device needed).  refs=(tid_lo, tid_hi): a shard."""
    L = load_library()
    p = dict(DEFAULT_PARAMS); p.update(params or {})
    P = Params(**p)
    lo, hi = (-1, -1) if refs is None else (int(refs[0]), int(refs[1]))
    flags = 2 if owns_unplaced else 0
    n = C.c_int64(0)
    _check(L.svtk_vcf_windows(os.fsencode(vcf_path), C.byref(P), int(n_ref), lo, hi, flags, 0, None, None, None, None, None, C.byref(n)))
    k = int(n.value)
    a = dict(chrom=np.zeros(k, np.int32), start=np.zeros(k, np.uint32), end=np.zeros(k, np.uint32), pos=np.zeros(k, np.uint32), kind=np.zeros(k, np.uint8))
    if k:
        _check(L.svtk_vcf_windows(os.fsencode(vcf_path), C.byref(P), int(n_ref), lo, hi, flags, k,
                                  *[a[x].ctypes.data for x in ("chrom", "start", "end", "pos", "kind")], C.byref(n)))
    return a


def vcf_format(vcf_path, results, out_path, params=None, n_ref=0, refs=None, owns_unplaced=True, info_lines=True):
    """svtk_vcf_format: the lines process_vcf writes for the given per-window results (no device needed); returns the record count."""
    L = load_library()
    p = dict(DEFAULT_PARAMS); p.update(params or {})
    P = Params(**p)
    lo, hi = (-1, -1) if refs is None else (int(refs[0]), int(refs[1]))
    flags = (2 if owns_unplaced else 0) | (0 if info_lines else 1)
    r = np.ascontiguousarray(results, dtype=np.int32)
    n = C.c_int64(0)
    _check(L.svtk_vcf_format(os.fsencode(vcf_path), C.byref(P), int(n_ref), lo, hi, flags, r.ctypes.data if r.size else None, int(r.size),
                             os.fsencode(out_path), C.byref(n)))
    return int(n.value)


class Auditor:
    """One GPU's refinement context (svtk_ctx): holds the decoded reads in HBM and runs batches of windows.

    Plays the role of one reference worker's t_arg (params.h:81-95): the tunables plus access to the reads."""

    def __init__(self, device=0, params=None, cuda_stream=None):
        self.lib = load_library()
        p = dict(DEFAULT_PARAMS); p.update(params or {})
        self._params = Params(**p)
        self.ctx = self.lib.svtk_create(int(device), C.byref(self._params), C.c_void_p(cuda_stream) if cuda_stream else None)
        if not self.ctx:
            raise SvtkError(f"svtk_create failed: {self.lib.svtk_last_error().decode()}")
        self._keep = []

    def close(self):
        if getattr(self, "ctx", None):
            self.lib.svtk_destroy(self.ctx)
            self.ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # ---- tunables (audt_args, params.h:57-62)
    def set_params(self, **kw):
        for k, v in kw.items():
            if k not in DEFAULT_PARAMS:
                raise KeyError(k)
            setattr(self._params, k, int(v))
        _check(self.lib.svtk_set_params(self.ctx, C.byref(self._params)))

    @property
    def params(self):
        return {k: getattr(self._params, k) for k in DEFAULT_PARAMS}

    # ---- reads
    def load_reads(self, rs):
        """Upload a host SoA read set (anything with tid_off/pos/end/n_cigar/cigar_off/cigar numpy arrays)."""
        arrs = [np.ascontiguousarray(rs.tid_off, dtype=np.int64), np.ascontiguousarray(rs.pos, dtype=np.uint32),
                np.ascontiguousarray(rs.end, dtype=np.uint32), np.ascontiguousarray(rs.n_cigar, dtype=np.uint32),
                np.ascontiguousarray(rs.cigar_off, dtype=np.uint64), np.ascontiguousarray(rs.cigar, dtype=np.uint32)]
        r = Reads(int(arrs[0].shape[0] - 1), int(arrs[1].shape[0]), int(arrs[5].shape[0]), *[a.ctypes.data for a in arrs])
        _check(self.lib.svtk_load_reads(self.ctx, C.byref(r)))

    def attach_reads_device(self, tid_off, pos_ptr, end_ptr, n_cigar_ptr, cigar_off_ptr, cigar_ptr, n_reads, n_words):
        """Adopt device-resident arrays (raw device pointers, e.g. torch tensor .data_ptr()); tid_off is a host array."""
        t = np.ascontiguousarray(tid_off, dtype=np.int64)
        r = Reads(int(t.shape[0] - 1), int(n_reads), int(n_words), t.ctypes.data, pos_ptr, end_ptr, n_cigar_ptr, cigar_off_ptr, cigar_ptr)
        _check(self.lib.svtk_attach_reads_device(self.ctx, C.byref(r)))

    # ---- batch seam
    def refine_windows(self, chrom, start, end, pos, kind):
        """N refine_start/refine_end/refine_ins/refine_point calls; host arrays in, int32 results out (-1 == NA)."""
        chrom = np.ascontiguousarray(chrom, dtype=np.int32)
        start, end, pos = _u32(start), _u32(end), _u32(pos)
        kind = np.ascontiguousarray(kind, dtype=np.uint8)
        n = int(chrom.shape[0])
        out = np.empty(n, dtype=np.int32)
        w = Windows(n, chrom.ctypes.data, start.ctypes.data, end.ctypes.data, pos.ctypes.data, kind.ctypes.data)
        _check(self.lib.svtk_refine_windows(self.ctx, C.byref(w), out.ctypes.data))
        return out

    def refine_windows_device(self, n, chrom_ptr, start_ptr, end_ptr, pos_ptr, kind_ptr, result_ptr):
        w = Windows(int(n), chrom_ptr, start_ptr, end_ptr, pos_ptr, kind_ptr)
        _check(self.lib.svtk_refine_windows_device(self.ctx, C.byref(w), result_ptr))

    def audit_device(self, n, chrom_ptr, start_ptr, end_ptr, pos_ptr, kind_ptr, result_ptr):
        """svtk_audit_device: index rebuild + every window, queued on the ctx stream; sync() finishes all queued passes."""
        w = Windows(int(n), chrom_ptr, start_ptr, end_ptr, pos_ptr, kind_ptr)
        _check(self.lib.svtk_audit_device(self.ctx, C.byref(w), result_ptr))

    def measure_windows(self, n, chrom_ptr, start_ptr, end_ptr, pos_ptr, kind_ptr, result_ptr):
        w = Windows(int(n), chrom_ptr, start_ptr, end_ptr, pos_ptr, kind_ptr)
        st = Stats()
        _check(self.lib.svtk_measure_windows(self.ctx, C.byref(w), result_ptr, C.byref(st)))
        return {k: int(getattr(st, k)) for k, _ in Stats._fields_}

    def measure_index(self, n, chrom_ptr, start_ptr, end_ptr, pos_ptr, kind_ptr, result_ptr):
        w = Windows(int(n), chrom_ptr, start_ptr, end_ptr, pos_ptr, kind_ptr)
        st = IndexStats()
        _check(self.lib.svtk_measure_index(self.ctx, C.byref(w), result_ptr, C.byref(st)))
        return {k: int(getattr(st, k)) for k, _ in IndexStats._fields_}

    def index_reads(self):
        """(Re)build the signature index of the resident reads (one pass over every CIGAR word), asynchronously."""
        _check(self.lib.svtk_index_reads(self.ctx))

    def sync(self):
        _check(self.lib.svtk_sync(self.ctx))

    @property
    def launch_count(self):
        return int(self.lib.svtk_launch_count(self.ctx))

    def set_profiling(self, on=True):
        """True / 1: CUDA events around every index / refine kernel launch; n > 1: around every n-th launch of each; False: off."""
        _check(self.lib.svtk_set_profiling(self.ctx, int(on)))

    def timing(self):
        t = Timing()
        _check(self.lib.svtk_get_timing(self.ctx, C.byref(t)))
        return {k: getattr(t, k) for k, _ in Timing._fields_}

    # ---- BAM in (svtk_load_bam): what hts_open + sam_index_load + the per-window iterators give the reference
    def load_bam(self, bam_path, bai_path=None, refs=None):
        """BGZF inflate + BAM record extraction on the GPU; returns the loader's counters / timings.
        refs=(tid_lo, tid_hi): only the reads of those references (a region shard), only their part of the file."""
        st = BamStats()
        bai = os.fsencode(bai_path) if bai_path else None
        if refs is None:
            _check(self.lib.svtk_load_bam(self.ctx, os.fsencode(bam_path), bai, C.byref(st)))
        else:
            _check(self.lib.svtk_load_bam_refs(self.ctx, os.fsencode(bam_path), bai, int(refs[0]), int(refs[1]), C.byref(st)))
        return {k: getattr(st, k) for k, _ in BamStats._fields_}

    def load_bam_windows(self, bam_path, chrom, start, end, bai_path=None):
        """svtk_load_bam_windows: only the parts of the file the given windows' iterators could touch."""
        chrom = np.ascontiguousarray(chrom, dtype=np.int32)
        start, end = _u32(start), _u32(end)
        w = Windows(int(chrom.shape[0]), chrom.ctypes.data, start.ctypes.data, end.ctypes.data, None, None)
        st = BamStats()
        _check(self.lib.svtk_load_bam_windows(self.ctx, os.fsencode(bam_path), os.fsencode(bai_path) if bai_path else None, C.byref(w), C.byref(st)))
        return {k: getattr(st, k) for k, _ in BamStats._fields_}

    def bam_refs(self):
        """[(name, length)] of the BAM header's reference dictionary."""
        out = []
        for t in range(self.lib.svtk_bam_n_ref(self.ctx)):
            nm, ln = C.c_char_p(), C.c_uint32()
            _check(self.lib.svtk_bam_ref(self.ctx, t, C.byref(nm), C.byref(ln)))
            out.append((nm.value.decode(), int(ln.value)))
        return out

    def ins_sequences(self, chrom, start, end, pos, kind, refined, tol=5, min_len=50):
        """svtk_ins_sequences: per INS window with a refined position, the inserted bases of the `I >= min_len` ops within
        `tol` of it, as read sets in poa_consensus' layout (copies of the ctx-owned arrays)."""
        chrom = np.ascontiguousarray(chrom, dtype=np.int32)
        start, end, pos = _u32(start), _u32(end), _u32(pos)
        kind = np.ascontiguousarray(kind, dtype=np.uint8)
        refined = np.ascontiguousarray(refined, dtype=np.int32)
        n = int(chrom.shape[0])
        w = Windows(n, chrom.ctypes.data, start.ctypes.data, end.ctypes.data, pos.ctypes.data, kind.ctypes.data)
        o = InsSets()
        _check(self.lib.svtk_ins_sequences(self.ctx, C.byref(w), refined.ctypes.data, int(tol), int(min_len), C.byref(o)))
        ns, nq = int(o.n_sets), int(o.n_seqs)

        def arr(ptr, count, dt):
            return np.ctypeslib.as_array(ptr, shape=(count,)).astype(dt, copy=True) if count else np.zeros(0, dt)
        seq_off = arr(o.seq_off, nq + 1, np.int64) if nq else np.zeros(1, np.int64)
        return dict(set_off=arr(o.set_off, ns + 1, np.int64) if ns else np.zeros(1, np.int64), seq_off=seq_off,
                    bases=arr(o.bases, int(seq_off[-1]), np.uint8), window=arr(o.window, ns, np.int32), read=arr(o.read, nq, np.int32),
                    ref_pos=arr(o.ref_pos, nq, np.uint32))

    def copy_reads(self):
        """The resident read set as host arrays (dict with the svtk_reads field names)."""
        nt, nr, nw = C.c_int64(), C.c_int64(), C.c_int64()
        _check(self.lib.svtk_copy_reads(self.ctx, C.byref(nt), C.byref(nr), C.byref(nw), None, None, None, None, None, None))
        a = dict(tid_off=np.zeros(nt.value + 1, np.int64), pos=np.zeros(nr.value, np.uint32), end=np.zeros(nr.value, np.uint32),
                 n_cigar=np.zeros(nr.value, np.uint32), cigar_off=np.zeros(nr.value, np.uint64), cigar=np.zeros(nw.value, np.uint32))
        _check(self.lib.svtk_copy_reads(self.ctx, None, None, None, *[a[k].ctypes.data for k in ("tid_off", "pos", "end", "n_cigar", "cigar_off", "cigar")]))
        return a

    def load_reads_ptr(self, n_tid, n_reads, n_words, tid_off_ptr, pos_ptr, end_ptr, n_cigar_ptr, cigar_off_ptr, cigar_ptr):
        """svtk_load_reads on raw HOST pointers (e.g. pinned torch tensors): no numpy conversion on the timed path."""
        r = Reads(int(n_tid), int(n_reads), int(n_words), tid_off_ptr, pos_ptr, end_ptr, n_cigar_ptr, cigar_off_ptr, cigar_ptr)
        _check(self.lib.svtk_load_reads(self.ctx, C.byref(r)))

    def refine_windows_ptr(self, n, chrom_ptr, start_ptr, end_ptr, pos_ptr, kind_ptr, result_ptr):
        """svtk_refine_windows on raw HOST pointers."""
        w = Windows(int(n), chrom_ptr, start_ptr, end_ptr, pos_ptr, kind_ptr)
        _check(self.lib.svtk_refine_windows(self.ctx, C.byref(w), result_ptr))

    # ---- per-record seam (refinement.h:25,42,60)
    def deletion(self, chrom, begin, end, sv_inter):
        res = Interval()
        _check(self.lib.svtk_deletion(self.ctx, int(chrom), Interval(*[x & 0xFFFFFFFF for x in begin]), Interval(*[x & 0xFFFFFFFF for x in end]),
                                      Interval(*[x & 0xFFFFFFFF for x in sv_inter]), C.byref(res)))
        return res.start, res.end

    def insertion(self, chrom, begin, pos):
        res = C.c_uint32()
        _check(self.lib.svtk_insertion(self.ctx, int(chrom), Interval(*[x & 0xFFFFFFFF for x in begin]), pos & 0xFFFFFFFF, C.byref(res)))
        return res.value

    def inversion(self, chrom, begin, end, sv_inter):
        res = Interval()
        _check(self.lib.svtk_inversion(self.ctx, int(chrom), Interval(*[x & 0xFFFFFFFF for x in begin]), Interval(*[x & 0xFFFFFFFF for x in end]),
                                       Interval(*[x & 0xFFFFFFFF for x in sv_inter]), C.byref(res)))
        return res.start, res.end

    # ---- pieces (parity aids)
    def consensus_pos_batch(self, values, offsets, pos):
        values = np.ascontiguousarray(values, dtype=np.int32)
        offsets = np.ascontiguousarray(offsets, dtype=np.int64)
        pos = np.ascontiguousarray(pos, dtype=np.int32)
        n = int(pos.shape[0])
        out = np.empty(n, dtype=np.int32)
        _check(self.lib.svtk_consensus_pos_batch(self.ctx, values.ctypes.data, offsets.ctypes.data, n, pos.ctypes.data, out.ctypes.data))
        return out

    def consensus_pos_batch_static(self, dev_values, dev_offsets, n_seg, n_values, dev_pos, dev_result):
        """svtk_consensus_pos_batch_static: raw DEVICE pointers in and out, no allocation, asynchronous on the ctx stream
        (call sync() before reading dev_result).  No counterpart in the reference (include/svtk.h)."""
        _check(self.lib.svtk_consensus_pos_batch_static(self.ctx, dev_values, dev_offsets, int(n_seg), int(n_values), dev_pos, dev_result))

    def consensus_pos(self, locations, pos):
        """consensus_pos(locations, size, pos, ...) of refinement.c:41 with the ctx's min-count/interval/range."""
        v = np.ascontiguousarray(locations, dtype=np.int32)
        return int(self.consensus_pos_batch(v, [0, v.shape[0]], [pos])[0])

    def window_signatures(self, kind, chrom, start, end, cap=1 << 16):
        buf = np.empty(cap, dtype=np.int32)
        n = C.c_int64(0)
        _check(self.lib.svtk_window_signatures(self.ctx, int(kind), int(chrom), Interval(start & 0xFFFFFFFF, end & 0xFFFFFFFF), buf.ctypes.data, cap, C.byref(n)))
        if n.value > cap:
            return self.window_signatures(kind, chrom, start, end, cap=int(n.value))
        return buf[:n.value].copy()

    def sliding_window_ins(self, chrom, start, end, window_size, slide_size, want_lines=False):
        best = C.c_int32(-1)
        cap = 1 << 22
        buf = C.create_string_buffer(cap) if want_lines else None
        _check(self.lib.svtk_sliding_window_ins(self.ctx, int(chrom), Interval(start & 0xFFFFFFFF, end & 0xFFFFFFFF), int(window_size), int(slide_size),
                                                C.byref(best), buf, cap if want_lines else 0))
        return (best.value, buf.value.decode()) if want_lines else best.value

    # ---- K3: insertion consensus (abpoa_msa, abPOA/src/abpoa_align.c:401)
    def poa_consensus(self, set_off, seq_off, bases, cons_cap=None, want_cov=False, want_trace=False):
        """abpoa_msa (default parameters) on every read set of the batch.  bases: uint8 codes A0 C1 G2 T3 N4.
        Returns (list of consensus code arrays, stats dict[, coverage list][, trace dict of per-read arrays])."""
        set_off = np.ascontiguousarray(set_off, dtype=np.int64)
        seq_off = np.ascontiguousarray(seq_off, dtype=np.int64)
        bases = np.ascontiguousarray(bases, dtype=np.uint8)
        ns = int(set_off.shape[0] - 1)
        n_reads = int(seq_off.shape[0] - 1)
        if cons_cap is None:                                   # room per set: twice its longest read + 64
            lens = np.diff(seq_off)
            cons_cap = np.array([2 * int(lens[set_off[s]:set_off[s + 1]].max(initial=0)) + 64 for s in range(ns)], dtype=np.int64)
        cons_off = np.zeros(ns + 1, dtype=np.int64)
        cons_off[1:] = np.cumsum(np.asarray(cons_cap, dtype=np.int64))
        cons = np.zeros(max(int(cons_off[-1]), 1), dtype=np.uint8)
        cov = np.zeros(max(int(cons_off[-1]), 1), dtype=np.int32) if want_cov else None
        clen = np.zeros(max(ns, 1), dtype=np.int32)
        status = np.zeros(max(ns, 1), dtype=np.int32)
        b = PoaBatch(ns, set_off.ctypes.data, seq_off.ctypes.data, bases.ctypes.data)
        st = PoaStats()
        tr = None
        tra = {}
        if want_trace:
            tra = dict(best_score=np.zeros(max(n_reads, 1), np.int32), node_n=np.zeros(max(n_reads, 1), np.int32), n_ops=np.zeros(max(n_reads, 1), np.int32),
                       ops_hash=np.zeros(max(n_reads, 1), np.uint64), cells=np.zeros(max(n_reads, 1), np.int64))
            tr = PoaTrace(*[tra[k].ctypes.data for k in ("best_score", "node_n", "n_ops", "ops_hash", "cells")])
        _check(self.lib.svtk_poa_consensus(self.ctx, C.byref(b), cons.ctypes.data, cons_off.ctypes.data, clen.ctypes.data,
                                           cov.ctypes.data if want_cov else None, status.ctypes.data, C.byref(tr) if tr else None, C.byref(st)))
        if ns and (clen[:ns] > np.diff(cons_off)).any():       # a consensus longer than its room: rerun with the reported lengths
            return self.poa_consensus(set_off, seq_off, bases, cons_cap=np.maximum(clen[:ns], 1) + 8, want_cov=want_cov, want_trace=want_trace)
        out = [cons[cons_off[s]:cons_off[s] + clen[s]].copy() for s in range(ns)]
        res = [out, {k: (list(getattr(st, k)) if k == "phase_cycles" else getattr(st, k)) for k, _ in PoaStats._fields_}]
        if want_cov:
            res.append([cov[cons_off[s]:cons_off[s] + clen[s]].copy() for s in range(ns)])
        if want_trace:
            res.append({k: v[:n_reads] for k, v in tra.items()})
        return tuple(res)

    # ---- whole path (audit.c:250 process_vcf)
    def process_vcf_ins(self, vcf_path, out_path, fasta_path, tol=-1):
        """process_vcf + a FASTA side file with the POA consensus of every refined INS record; returns (records, consensuses)."""
        n, k = C.c_int64(0), C.c_int64(0)
        _check(self.lib.svtk_process_vcf_ins(self.ctx, vcf_path.encode(), out_path.encode() if out_path else None, fasta_path.encode(), int(tol),
                                             C.byref(n), C.byref(k)))
        return n.value, k.value

    def process_vcf_region(self, vcf_path, out_path, tid_lo, tid_hi, info_lines=False, owns_unplaced=False):
        """svtk_process_vcf_region: the records on references tid_lo..tid_hi only (one GPU's region shard)."""
        n = C.c_int64(0)
        flags = (0 if info_lines else 1) | (2 if owns_unplaced else 0)
        _check(self.lib.svtk_process_vcf_region(self.ctx, vcf_path.encode(), out_path.encode() if out_path else None, int(tid_lo), int(tid_hi), flags, C.byref(n)))
        return n.value

    def audit_shard(self, bam_path, vcf_path, out_path, shard, n_shards, info_lines=False, bai_path=None):
        """svtk_audit_shard: region shard `shard` of n_shards of the whole job, file to file; returns (records, loader stats)."""
        n, st = C.c_int64(0), BamStats()
        _check(self.lib.svtk_audit_shard(self.ctx, os.fsencode(bam_path), os.fsencode(bai_path) if bai_path else None, os.fsencode(vcf_path),
                                         os.fsencode(out_path), int(shard), int(n_shards), 0 if info_lines else 1, C.byref(n), C.byref(st)))
        return n.value, {k: getattr(st, k) for k, _ in BamStats._fields_}

    def process_vcf(self, vcf_path, out_path):
        n = C.c_int64(0)
        _check(self.lib.svtk_process_vcf(self.ctx, vcf_path.encode(), out_path.encode() if out_path else None, C.byref(n)))
        return n.value
