// svtk_index.cu -- the streaming form of the audt refinement path (sm_100a).
//
// The reference walks a read's CIGAR once per window that sees it (refinement.c:117-160 inside every refine_*
// call), i.e. every CIGAR word is decoded ~7 times at 30x.  What a window can take from a read depends on the
// window only through two comparisons:
//     - a D/I op is emitted iff the reference cursor BEFORE it has not passed W.end   (stop rule, :141-144)
//     - the clip rules compare the read's start / final cursor with [W.start, W.end]  (:147-159, :210-221)
// so the walk itself is window independent.  This file splits the path accordingly:
//
//   k_index_reads     ONE coalesced pass over all CIGAR words (HBM-bound, 128-bit loads, block-wide scan of the
//                     reference cursor): every read yields its signature events {cursor, len|type, read, endpos}
//                     -- D > 50, I >= 50, trailing / leading soft clip -- and its final cursor.  Events of a
//                     chunk of 32 reads land contiguously in HBM (one atomicAdd per chunk).
//   k_refine_indexed  one CTA per window: 32-ary search of the read range the BAM iterator would yield
//                     (hts.c:3049-3090, 3905-3921), gather of the events of the covering chunks under the two
//                     rules above into shared memory, bitonic sort, consensus_pos (refinement.c:41-101).
//
// Results are bit-identical to the per-(window, read) walker in svtk_kernels.cu (kept as the instrumented
// cross-check behind svtk_measure_windows).
#include "svtk_device.cuh"
#include "svtk_launch.h"

namespace svtk {

constexpr int EV_DEL = 0, EV_INS = 1, EV_TCLIP = 2, EV_LCLIP = 3;

constexpr int IDX_NT = 256;                 // threads per CTA
constexpr int IDX_WPT = 8;                  // CIGAR words per thread and tile
constexpr int IDX_TILE = IDX_NT * IDX_WPT;  // 2048 words = 8 KB per tile
constexpr int IDX_EVCAP = 512;              // events staged in shared memory between flushes

struct DevIndex {
    uint4 *events;       // {x, len << 4 | type, read index, read endpos}
    uint4 *chunks;       // per chunk of IDX_READS reads: newest segment {base, count, next segment or -1, 0}
    uint4 *segs;         // older segments of chunks that flushed more than once
    uint32_t *rend;      // per read: reference cursor after the whole CIGAR
    unsigned int *ctrl;  // [0] event top  [1] segment top  [2] error bits (1 event pool, 2 segment pool)
    unsigned int ev_cap, seg_cap;
    int n_chunks;
};

static DevIndex to_dev(const LaunchIndex &x) {
    DevIndex d;
    d.events = reinterpret_cast<uint4 *>(x.events); d.chunks = reinterpret_cast<uint4 *>(x.chunks);
    d.segs = reinterpret_cast<uint4 *>(x.segs); d.rend = x.rend; d.ctrl = x.ctrl; d.ev_cap = x.ev_cap; d.seg_cap = x.seg_cap;
    d.n_chunks = x.n_chunks;
    return d;
}
static DevReads to_dev(const LaunchReads &r) {
    DevReads d;
    d.n_tid = r.n_tid; d.n_reads = r.n_reads; d.tid_off = r.tid_off; d.pos = r.pos; d.end = r.end; d.n_cigar = r.n_cigar;
    d.cigar_off = r.cigar_off; d.cigar = r.cigar; d.max_span = r.max_span;
    return d;
}
static DevWindows to_dev(const LaunchWindows &w) {
    DevWindows d;
    d.n_win = w.n_win; d.chrom = w.chrom; d.start = w.start; d.end = w.end; d.pos = w.pos; d.kind = w.kind;
    return d;
}

// ---------------------------------------------------------------------------------------------------------
struct IndexShared {
    uint32_t P[IDX_TILE];                // exclusive cursor prefix (chunk relative) of every word of the tile
    unsigned short cand[IDX_TILE];       // tile-local indices of D > 50 / I >= 50 words
    uint4 ev[IDX_EVCAP];
    uint32_t rel[IDX_READS + 1];         // word offset of each read, relative to the aligned chunk start
    uint32_t pos[IDX_READS], endp[IDX_READS];
    uint32_t base[IDX_READS + 1];        // P at each read's first word
    uint32_t wtot[IDX_NT / 32];
    int ncand[2];
    int nclip;
    uint32_t flush_base;
    uint4 head;                          // newest segment of this chunk
};

// write the staged events as one more segment of the chunk (uniform call; nev > 0)
__device__ __forceinline__ void index_flush(IndexShared &s, const DevIndex &X, int nev) {
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned int b = atomicAdd(&X.ctrl[0], (unsigned int)nev);
        uint32_t fb = 0xffffffffu;
        if ((unsigned long long)b + (unsigned long long)nev > (unsigned long long)X.ev_cap) atomicOr(&X.ctrl[2], 1u);
        else {
            int next = -1;
            bool ok = true;
            if (s.head.y != 0u) {                            // push the previous segment down the chain
                const unsigned int si = atomicAdd(&X.ctrl[1], 1u);
                if (si >= X.seg_cap) { atomicOr(&X.ctrl[2], 2u); ok = false; }
                else { X.segs[si] = s.head; next = (int)si; }
            }
            if (ok) { s.head = make_uint4(b, (uint32_t)nev, (uint32_t)next, 0u); fb = b; }
        }
        s.flush_base = fb;
    }
    __syncthreads();
    const uint32_t fb = s.flush_base;
    if (fb != 0xffffffffu)
        for (int i = threadIdx.x; i < nev; i += IDX_NT) X.events[(size_t)fb + i] = s.ev[i];
    __syncthreads();
}

__global__ void __launch_bounds__(IDX_NT) k_index_reads(DevReads R, DevIndex X) {
    __shared__ IndexShared s;
    constexpr int NW = IDX_NT / 32;
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const long long r0 = (long long)blockIdx.x * IDX_READS;
    const int nr = (int)min((long long)IDX_READS, R.n_reads - r0);
    const unsigned long long w0 = __ldg(R.cigar_off + r0);
    const unsigned long long w0a = w0 & ~3ull;
    if (t <= nr) {
        const long long r = r0 + t;
        const unsigned long long off = (t < nr) ? __ldg(R.cigar_off + r) : __ldg(R.cigar_off + r - 1) + __ldg(R.n_cigar + r - 1);
        s.rel[t] = (uint32_t)(off - w0a);
        if (t < nr) { s.pos[t] = __ldg(R.pos + r); s.endp[t] = __ldg(R.end + r); }
    }
    if (t == 0) { s.ncand[0] = 0; s.ncand[1] = 0; s.nclip = 0; s.head = make_uint4(0u, 0u, 0xffffffffu, 0u); }
    __syncthreads();
    const uint32_t head_pad = (uint32_t)(w0 - w0a);
    const uint32_t span = s.rel[nr];
    const uint32_t *__restrict__ cg = R.cigar + w0a;

    uint32_t run = 0;       // cursor advance of all words before the current tile (uniform)
    int nev = 0;            // staged events (uniform)
    int par = 0;
    for (uint32_t tb = 0; tb < span; tb += IDX_TILE, par ^= 1) {
        const uint32_t i0 = tb + (uint32_t)t * IDX_WPT;
        uint32_t w[IDX_WPT];
#pragma unroll
        for (int q = 0; q < IDX_WPT / 4; ++q) {
            uint4 v = make_uint4(0u, 0u, 0u, 0u);
            if (i0 + 4u * q < span) v = __ldg(reinterpret_cast<const uint4 *>(cg + i0 + 4u * q));
            w[4 * q] = v.x; w[4 * q + 1] = v.y; w[4 * q + 2] = v.z; w[4 * q + 3] = v.w;
        }
        uint32_t p[IDX_WPT];
        uint32_t sum = 0;
        unsigned candm = 0;
#pragma unroll
        for (int k = 0; k < IDX_WPT; ++k) {
            const uint32_t gi = i0 + k;
            if (gi < head_pad || gi >= span) w[k] = 0u;      // the neighbouring chunks' words: "0M", inert
            const uint32_t op = w[k] & 15u, len = w[k] >> 4;
            p[k] = sum;
            sum += ((0x12u >> op) & 1u) ? 0u : len;          // every op except I(1) and S(4) advances (refinement.c:137)
            if ((op == (uint32_t)OP_DEL && len > (uint32_t)SV_MIN_LENGTH) || (op == (uint32_t)OP_INS && len >= (uint32_t)SV_MIN_LENGTH))
                candm |= 1u << k;
        }
        const uint32_t incl = warp_inclusive_scan(sum);
        if (lane == 31) s.wtot[warp] = incl;
        __syncthreads();                                     // #1
        if (t == 0) s.ncand[par ^ 1] = 0;
        uint32_t woff = 0, ttot = 0;
#pragma unroll
        for (int v = 0; v < NW; ++v) { const uint32_t x = s.wtot[v]; ttot += x; if (v < warp) woff += x; }
        const uint32_t mybase = run + woff + (incl - sum);
#pragma unroll
        for (int q = 0; q < IDX_WPT / 4; ++q)
            *reinterpret_cast<uint4 *>(&s.P[t * IDX_WPT + 4 * q]) =
                make_uint4(mybase + p[4 * q], mybase + p[4 * q + 1], mybase + p[4 * q + 2], mybase + p[4 * q + 3]);
        if (candm) {
            const int slot = atomicAdd(&s.ncand[par], __popc(candm));
            int j = 0;
#pragma unroll
            for (int k = 0; k < IDX_WPT; ++k)
                if (candm & (1u << k)) s.cand[slot + j++] = (unsigned short)(t * IDX_WPT + k);
        }
        __syncthreads();                                     // #2
        if (t < nr) {
            const uint32_t rel = s.rel[t];
            if (rel >= tb && rel < tb + IDX_TILE && rel < span) s.base[t] = s.P[rel - tb];
        }
        __syncthreads();                                     // #3
        const int ncand = s.ncand[par];
        int done = 0;
        while (done < ncand) {
            if (nev == IDX_EVCAP) { index_flush(s, X, nev); nev = 0; }
            const int take = min(IDX_EVCAP - nev, ncand - done);
            for (int ci = t; ci < take; ci += IDX_NT) {
                const uint32_t i = s.cand[done + ci];
                const uint32_t gi = tb + i;
                int lo = 0, hi = nr;                         // the read owning word gi: last r with rel[r] <= gi
                while (hi - lo > 1) { const int m = (lo + hi) >> 1; if (s.rel[m] <= gi) lo = m; else hi = m; }
                const uint32_t word = __ldg(cg + gi);
                const uint32_t x = s.pos[lo] + (s.P[i] - s.base[lo]);
                const uint32_t type = ((word & 15u) == (uint32_t)OP_DEL) ? EV_DEL : EV_INS;
                s.ev[nev + ci] = make_uint4(x, (word & ~15u) | type, (uint32_t)(r0 + lo), s.endp[lo]);
            }
            nev += take; done += take;
        }
        run += ttot;
    }
    // reads that start at the very end of the chunk's span (empty CIGARs) and the end-of-span sentinel
    if (t <= nr && s.rel[t] >= span) s.base[t] = run;
    __syncthreads();
    if (nev + 2 * IDX_READS > IDX_EVCAP && nev > 0) { index_flush(s, X, nev); nev = 0; }
    if (t < nr) {
        const uint32_t n = s.rel[t + 1] - s.rel[t];
        uint32_t rendv = s.pos[t];
        if (n > 0) {
            rendv += s.base[t + 1] - s.base[t];
            const uint32_t first = __ldg(cg + s.rel[t]), last = __ldg(cg + s.rel[t + 1] - 1);
            if ((last & 15u) == (uint32_t)OP_SOFT) {         // refinement.c:120,147: trailing soft clip, value = final cursor
                const int k = atomicAdd(&s.nclip, 1);
                s.ev[nev + k] = make_uint4(rendv, EV_TCLIP, (uint32_t)(r0 + t), s.endp[t]);
            }
            if ((first & 15u) == (uint32_t)OP_SOFT) {        // refinement.c:210: leading soft clip, keyed by core.pos
                const int k = atomicAdd(&s.nclip, 1);
                s.ev[nev + k] = make_uint4(s.pos[t], EV_LCLIP, (uint32_t)(r0 + t), s.endp[t]);
            }
        }
        X.rend[r0 + t] = rendv;
    }
    __syncthreads();
    nev += s.nclip;
    if (nev > 0) index_flush(s, X, nev);
    if (t == 0) X.chunks[blockIdx.x] = s.head;
}

// ---------------------------------------------------------------------------------------------------------
// first index in [a,b) with pos[idx] >= key, by the whole warp: 32 probes per round instead of one
__device__ __forceinline__ long long warp_lower_bound_pos(const uint32_t *__restrict__ pos, long long a, long long b, long long key) {
    const int lane = lane_id();
    while (b - a > 32) {
        const long long step = (b - a + 31) / 32;
        long long idx = a + (long long)(lane + 1) * step - 1;
        if (idx > b - 1) idx = b - 1;
        const bool ge = (long long)__ldg(pos + idx) >= key;
        const unsigned m = __ballot_sync(FULL, ge);
        if (!m) return b;
        const int L = __ffs(m) - 1;
        const long long na = a + (long long)L * step;
        b = __shfl_sync(FULL, idx, L);                       // pos[b] >= key: the answer lies in [na, b]
        a = na;
    }
    const long long idx = a + lane;
    const bool ge = idx < b && (long long)__ldg(pos + idx) >= key;
    const unsigned m = __ballot_sync(FULL, ge);
    return m ? a + (__ffs(m) - 1) : b;
}

// cursor right after the op that first pushes it past w_end (refinement.c:141-144); the caller knows it exists
__device__ __forceinline__ uint32_t warp_stop_cursor(const uint32_t *__restrict__ cigar, unsigned long long off, uint32_t n,
                                                     uint32_t rpos, uint32_t w_end) {
    const int lane = lane_id();
    uint32_t carry = rpos;
    for (uint32_t b = 0; b < n; b += 32) {
        const uint32_t w = (b + lane < n) ? __ldg(cigar + off + b + lane) : 0u;
        const uint32_t op = w & 15u;
        const uint32_t adv = ((0x12u >> op) & 1u) ? 0u : (w >> 4);
        const uint32_t incl = warp_inclusive_scan(adv);
        const uint32_t after = carry + incl;
        const unsigned m = __ballot_sync(FULL, after > w_end);
        if (m) return __shfl_sync(FULL, after, __ffs(m) - 1);
        carry += __shfl_sync(FULL, incl, 31);
    }
    return carry;
}

struct RefineShared {
    long long lo, hi;
    int count;
    int n_defer;
    int defer[32];                    // reads whose leading-clip rule needs the stop cursor (rare)
    ScanResult res[2];
};

struct IndexCounters {                // instrumented instantiation only
    unsigned long long events_scanned, segments, pairs_range, signatures, overflow;
};

// One CTA per window.  NT threads, CAP-int signature segment in shared memory; windows with more signatures go to
// the overflow list exactly like the fused walker's.
template <int NT, int CAP, bool STATS>
__global__ void __launch_bounds__(NT) k_refine_indexed(DevReads R, DevWindows W, DevIndex X, Tunables t, int32_t *__restrict__ result,
                                                       OverflowList ov, int2 *__restrict__ range_out, IndexCounters *ic) {
    __shared__ int s_sig[CAP];
    __shared__ RefineShared s;
    constexpr int NWARP = NT / 32;
    const int w = blockIdx.x;
    const int warp = threadIdx.x >> 5, lane = lane_id();
    const int kind = W.kind[w];
    const uint32_t w_start = W.start[w], w_end = W.end[w];
    const int tid = W.chrom[w] - 1;                              // refinement.c:114
    const uint32_t beg = w_start - 1u;
    const long long begq = (long long)beg, endq = (long long)(uint32_t)(w_end - 1u);   // uint32 arithmetic, then widened
    const bool live = tid >= 0 && tid < R.n_tid && endq >= begq && kind != KIND_POINT;
    if (threadIdx.x == 0) { s.count = 0; s.n_defer = 0; s.lo = 0; s.hi = 0; }
    __syncthreads();
    if (live && warp < 2) {
        const long long t0 = R.tid_off[tid], t1 = R.tid_off[tid + 1];
        // yielded reads have pos < endq (hts.c:3913) and pos + span > beg; span <= max_span bounds the look-back
        const long long key = warp == 0 ? endq : begq - (long long)R.max_span + 1;
        const long long v = warp_lower_bound_pos(R.pos, t0, t1, key);
        if (lane == 0) { if (warp == 0) s.hi = v; else s.lo = v; }
    }
    __syncthreads();
    const long long hi = s.hi, lo = s.lo < s.hi ? s.lo : s.hi;
    if (range_out && threadIdx.x == 0) range_out[w] = make_int2((int)lo, (int)hi);
    if (hi <= lo) {                                              // no iterator / no reads / refine_point
        if (threadIdx.x == 0) result[w] = -1;
        return;
    }
    const SigSink sink{s_sig, &s.count, CAP};
    const long long c0 = lo / IDX_READS, c1 = (hi - 1) / IDX_READS;
    unsigned long long n_scan = 0, n_seg = 0;
    for (long long c = c0 + warp; c <= c1; c += NWARP) {
        uint4 seg = __ldg(X.chunks + c);
        while (true) {
            const uint32_t base = seg.x, cnt = seg.y;
            if (STATS) { n_scan += cnt; n_seg += 1; }
            for (uint32_t e0 = 0; e0 < cnt; e0 += 32) {
                const uint32_t e = e0 + lane;
                bool emit = false;
                int val = 0;
                if (e < cnt) {
                    const uint4 ev = __ldg(X.events + (size_t)base + e);
                    const uint32_t type = ev.y & 15u, len = ev.y >> 4;
                    const long long ri = (long long)ev.z;
                    if (ri >= lo && ri < hi && ev.w > beg) {         // the iterator yields this read (endpos > beg, hts.c:3914)
                        if (kind == KIND_DEL_START) {
                            if (type == EV_DEL) { emit = ev.x <= w_end; val = (int)ev.x; }
                            else if (type == EV_TCLIP) { emit = w_start <= ev.x && ev.x <= w_end; val = (int)ev.x; }
                        } else if (kind == KIND_DEL_END) {
                            if (type == EV_DEL) { emit = ev.x <= w_end; val = (int)(ev.x + len + 1u); }
                            else if (type == EV_LCLIP && w_start <= ev.x && ev.x <= w_end) {
                                const uint32_t rendv = __ldg(X.rend + ri);
                                if (rendv <= w_end) { emit = true; val = (int)(rendv + 1u); }    // walk never stopped: final cursor + 1
                                else { const int k = atomicAdd(&s.n_defer, 1); if (k < 32) s.defer[k] = (int)ri; }
                            }
                        } else {                                     // KIND_INS
                            if (type == EV_INS) { emit = ev.x <= w_end; val = (int)ev.x; }
                        }
                    }
                }
                const unsigned m = __ballot_sync(FULL, emit);
                if (m) {
                    int b = 0;
                    if (lane == 0) b = atomicAdd(&s.count, __popc(m));
                    b = __shfl_sync(FULL, b, 0);
                    if (emit) { const int slot = b + __popc(m & ((1u << lane) - 1u)); if (slot < CAP) s_sig[slot] = val; }
                }
            }
            if (seg.z == 0xffffffffu) break;
            seg = __ldg(X.segs + seg.z);
        }
    }
    __syncthreads();
    // leading-clip reads whose walk stops inside the window: value = cursor after the crossing op, + 1 (refinement.c:210-220)
    const int nd = s.n_defer;
    if (nd > 0) {
        if (nd <= 32) {
            for (int k = warp; k < nd; k += NWARP) {
                const long long ri = s.defer[k];
                const uint32_t sr = warp_stop_cursor(R.cigar, __ldg(R.cigar_off + ri), __ldg(R.n_cigar + ri), __ldg(R.pos + ri), w_end);
                if (lane == 0) sink.push((int)(sr + 1u));
            }
        } else {
            // more than the defer list holds: rescan the read range for them (still rare)
            for (long long ri = lo + warp; ri < hi; ri += NWARP) {
                const uint32_t n = __ldg(R.n_cigar + ri), rp = __ldg(R.pos + ri);
                if (n == 0 || !(__ldg(R.end + ri) > beg) || !(w_start <= rp && rp <= w_end)) continue;
                const unsigned long long off = __ldg(R.cigar_off + ri);
                if ((__ldg(R.cigar + off) & 15u) != (uint32_t)OP_SOFT || __ldg(X.rend + ri) <= w_end) continue;
                const uint32_t sr = warp_stop_cursor(R.cigar, off, n, rp, w_end);
                if (lane == 0) sink.push((int)(sr + 1u));
            }
        }
        __syncthreads();
    }
    const int n = s.count;
    if (STATS) {
        if (lane == 0) { atomicAdd(&ic->events_scanned, n_scan); atomicAdd(&ic->segments, n_seg); }
        if (threadIdx.x == 0) { atomicAdd(&ic->pairs_range, (unsigned long long)(hi - lo)); atomicAdd(&ic->signatures, (unsigned long long)n); }
    }
    if (n < t.min_count || n <= 0) {                             // refinement.c:43
        if (threadIdx.x == 0) result[w] = -1;
        return;
    }
    if (n > CAP) {
        if (threadIdx.x == 0) {
            const int slot = atomicAdd(ov.count, 1);
            if (slot < ov.capacity) ov.items[slot] = make_int2(w, n);
            result[w] = -1;
            if (STATS) atomicAdd(&ic->overflow, 1ull);
        }
        return;
    }
    block_bitonic_sort(s_sig, n);
    const int res = block_consensus(s_sig, n, (int)W.pos[w], t, s.res);
    if (threadIdx.x == 0) result[w] = res;
}

// =========================================================================================================
constexpr int RI_NT = 128;
constexpr int RI_CAP = 1024;

int index_reads_per_chunk() { return IDX_READS; }

cudaError_t launch_index_reads(const LaunchReads &r, const LaunchIndex &x, cudaStream_t s) {
    if (r.n_reads <= 0) return cudaSuccess;
    k_index_reads<<<x.n_chunks, IDX_NT, 0, s>>>(to_dev(r), to_dev(x));
    return cudaGetLastError();
}

cudaError_t launch_refine_indexed(const LaunchReads &r, const LaunchWindows &w, const LaunchIndex &x, int2 *range, int range_, int interval,
                                  int min_count, int32_t *result, int *ov_count, int2 *ov_items, int ov_capacity,
                                  unsigned long long *counters, cudaStream_t s) {
    if (w.n_win <= 0) return cudaSuccess;
    const Tunables t{range_, interval, min_count};
    const OverflowList ov{ov_count, ov_items, ov_capacity};
    if (counters)
        k_refine_indexed<RI_NT, RI_CAP, true><<<w.n_win, RI_NT, 0, s>>>(to_dev(r), to_dev(w), to_dev(x), t, result, ov, range,
                                                                         reinterpret_cast<IndexCounters *>(counters));
    else
        k_refine_indexed<RI_NT, RI_CAP, false><<<w.n_win, RI_NT, 0, s>>>(to_dev(r), to_dev(w), to_dev(x), t, result, ov, range, nullptr);
    return cudaGetLastError();
}

}  // namespace svtk

namespace svtk {

// svtk_reads contract check: CIGARs packed back to back in read order (cigar_off[r] + n_cigar[r] == cigar_off[r+1])
__global__ void k_check_layout(const unsigned long long *__restrict__ off, const uint32_t *__restrict__ n, long long n_reads,
                               unsigned long long n_words, unsigned int *flag) {
    for (long long r = blockIdx.x * (long long)blockDim.x + threadIdx.x; r < n_reads; r += (long long)gridDim.x * blockDim.x) {
        const unsigned long long next = (r + 1 < n_reads) ? off[r + 1] : n_words;
        if (off[r] + n[r] != next || (r == 0 && off[0] != 0ull)) atomicOr(flag, 1u);
    }
}

cudaError_t launch_check_layout(const LaunchReads &r, unsigned long long n_words, unsigned int *flag, cudaStream_t s) {
    if (r.n_reads <= 0) return cudaSuccess;
    const long long want = (r.n_reads + 255) / 256;
    const int blocks = (int)(want < 148 * 8 ? want : 148 * 8);
    k_check_layout<<<blocks, 256, 0, s>>>(r.cigar_off, r.n_cigar, r.n_reads, n_words, flag);
    return cudaGetLastError();
}

}  // namespace svtk
