// svtk_device.cuh -- device-side building blocks of the audt refinement path (sm_100a).
//
//   walk_read_warp     one warp walks one read's CIGAR: 128-bit coalesced loads of packed BAM CIGAR words,
//                      warp-scan of the reference cursor, stop rule, DEL/INS/soft-clip signature emission
//                      (reference: refinement.c:117-160, 183-222, 293-320)
//   block_bitonic_sort ascending sort of the per-window signature segment (reference: utils.c:4-25 -- any
//                      ascending sort of plain ints is bit-identical)
//   block_consensus    consensus_pos's two directional cluster scans with early return, one warp per
//                      direction (reference: refinement.c:41-101)
//   block_slide        sliding_window_ins's per-tile support scan (reference: sliding_window.c:65-92)
//
// All arithmetic is 32-bit integer with wrapping semantics, plus one 64-bit cluster sum, exactly as the
// reference computes it.  Nothing here is a dense contraction: no tensor cores, the roofline is HBM.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace svtk {

constexpr int KIND_DEL_START = 0, KIND_DEL_END = 1, KIND_INS = 2, KIND_POINT = 3;
constexpr int OP_INS = 1, OP_DEL = 2, OP_SOFT = 4;
constexpr int SV_MIN_LENGTH = 50;          // params.h:33
constexpr unsigned FULL = 0xffffffffu;
constexpr int WARP_WORDS = 128;            // CIGAR words one warp consumes per step: 32 lanes x uint4

struct DevReads {
    int n_tid;
    long long n_reads;
    const long long *tid_off;    // device copy, [n_tid+1]
    const uint32_t *pos, *end, *n_cigar;
    const unsigned long long *cigar_off;
    const uint32_t *cigar;       // 16-byte aligned, readable up to the next 16-byte boundary past the last word
    uint32_t max_span;           // max over reads of end - pos
    const uint32_t *tiles;       // position tiles (svtk_index.cu k_pos_tiles): first read with pos >= tile start, per reference; nullptr: none
    const long long *tile_off;   // [n_tid + 1] first tile of each reference
};
constexpr int POS_TILE_SHIFT = 12;          // 4 096-base tiles: ~8 reads per tile at 30x / 15 kb

struct DevWindows {
    int n_win;
    const int32_t *chrom;
    const uint32_t *start, *end, *pos;
    const uint8_t *kind;
};

struct Tunables { int range, interval, min_count; };

#ifndef SVTK_IDX_READS
#define SVTK_IDX_READS 8
#endif
constexpr int IDX_READS = SVTK_IDX_READS;               // reads per chunk of the signature index: one warp's share (svtk_index.cu)
constexpr int IDX_POOLS = 16;              // event sub-pools (spreads the per-flush atomicAdd over 16 L2 lines)

struct OverflowList {
    int *count;        // number of overflowed windows
    int2 *items;       // (window index, signature count)
    int capacity;
};

struct WorkCounters {            // svtk_stats, accumulated with atomics by the instrumented instantiation only
    unsigned long long pairs, words, signatures, overflow;
};

__device__ __forceinline__ int iabs_wrap(int a) { return a < 0 ? -a : a; }   // refinement.h:7 on int
__device__ __forceinline__ int lane_id() { return threadIdx.x & 31; }

// Signature sink: a bounded array (shared or global) with a running count that keeps counting past cap,
// so an overflowing window knows how much room its retry needs.
struct SigSink {
    int *buf;
    int *count;
    int cap;
    __device__ __forceinline__ void push(int v) const {
        int i = atomicAdd(count, 1);
        if (i < cap) buf[i] = v;
    }
};

__device__ __forceinline__ uint32_t warp_inclusive_scan(uint32_t v) {
    const int lane = lane_id();
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        uint32_t t = __shfl_up_sync(FULL, v, d);
        if (lane >= d) v += t;
    }
    return v;
}

__device__ __forceinline__ uint4 ldg_cigar4(const uint32_t *cigar, unsigned long long word_index) {
    return __ldg(reinterpret_cast<const uint4 *>(cigar + word_index));
}

// One warp, one read.  rpos = bam1_core_t.pos, n = n_cigar (> 0), off = word offset of the CIGAR.
// Returns nothing; signatures go to `sink`.  `visited` (instrumented builds) receives the number of CIGAR
// words the reference's loop touches for this (window, read) pair.
template <bool STATS>
__device__ __forceinline__ void walk_read_warp(const uint32_t *__restrict__ cigar, int kind, uint32_t w_start, uint32_t w_end,
                                               uint32_t rpos, uint32_t n, unsigned long long off, const SigSink &sink,
                                               unsigned long long &visited) {
    const int lane = lane_id();
    const unsigned long long aligned = off & ~3ull;
    const int head = (int)(off - aligned);                    // words of the previous read sharing our first uint4
    const uint32_t span = n + (uint32_t)head;                 // aligned-relative end of this read's words
    const uint32_t n_chunks = (span + WARP_WORDS - 1) / WARP_WORDS;
    const uint32_t want_op = (kind == KIND_INS) ? OP_INS : OP_DEL;
    const uint32_t min_len = (kind == KIND_INS) ? SV_MIN_LENGTH : SV_MIN_LENGTH + 1;   // I: 50 <= len ; D: 50 < len
    const bool emits = (kind != KIND_POINT);

    uint32_t carry = rpos;       // reference cursor at the start of the current chunk
    bool stopped = false;
    uint32_t stop_r = 0;         // cursor right after the op that crossed w_end

    const uint32_t my0 = (uint32_t)lane * 4u;
    uint4 cur = make_uint4(0, 0, 0, 0);
    if (my0 < span) cur = ldg_cigar4(cigar, aligned + my0);
    for (uint32_t c = 0; c < n_chunks; ++c) {
        // prefetch the next chunk of this read before touching the current one
        uint4 nxt = make_uint4(0, 0, 0, 0);
        const uint32_t nb = (c + 1) * WARP_WORDS + my0;
        if (nb < span) nxt = ldg_cigar4(cigar, aligned + nb);

        const int base = (int)(c * WARP_WORDS + my0) - head;  // read-relative index of my first word
        uint32_t w[4] = {cur.x, cur.y, cur.z, cur.w};
        uint32_t adv[4];
        uint32_t sum = 0;
        bool cand = false;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int idx = base + k;
            if (idx < 0 || idx >= (int)n) w[k] = 0;           // neighbour read's words: 0 == "0M", inert
            const uint32_t op = w[k] & 15u, len = w[k] >> 4;
            adv[k] = ((0x12u >> op) & 1u) ? 0u : len;         // every op except I(1) and S(4) advances (refinement.c:137)
            sum += adv[k];
            cand |= (op == want_op) & (len >= min_len);
        }
        const uint32_t incl = warp_inclusive_scan(sum);
        const uint32_t r_before = carry + (incl - sum);
        const uint32_t r_after = carry + incl;
        // a word is processed iff the cursor before it has not yet exceeded w_end (refinement.c:141-144)
        if (emits && cand && r_before <= w_end) {
            uint32_t r = r_before;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const uint32_t op = w[k] & 15u, len = w[k] >> 4;
                if (r <= w_end && op == want_op && len >= min_len)
                    sink.push((int)(kind == KIND_DEL_END ? r + len + 1u : r));
                r += adv[k];
            }
        }
        const bool crosses = (r_before <= w_end) && (r_after > w_end);
        const unsigned cm = __ballot_sync(FULL, crosses);
        if (cm) {
            const int src = __ffs(cm) - 1;
            uint32_t r = r_before, sr = 0;
            int sk = 0;
            if (lane == src) {
                bool found = false;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    r += adv[k];
                    if (!found && r > w_end) { found = true; sr = r; sk = base + k; }
                }
            }
            stop_r = __shfl_sync(FULL, sr, src);
            if (STATS) visited += (unsigned long long)(__shfl_sync(FULL, sk, src) + 1);
            stopped = true;
            break;
        }
        carry += __shfl_sync(FULL, incl, 31);
        cur = nxt;
    }
    if (STATS && !stopped) visited += n;

    if (lane == 0) {
        if (kind == KIND_DEL_START) {
            // trailing soft clip inside the window, only if the walk ran off the end of the CIGAR (refinement.c:120,142,147)
            if (!stopped) {
                const uint32_t last = __ldg(cigar + off + n - 1);
                if ((last & 15u) == OP_SOFT && w_start <= carry && carry <= w_end) sink.push((int)carry);
            }
        } else if (kind == KIND_DEL_END) {
            // leading soft clip with core.pos inside the window records the POST-walk cursor + 1 (refinement.c:210-220)
            const uint32_t first = __ldg(cigar + off);
            if ((first & 15u) == OP_SOFT && w_start <= rpos && rpos <= w_end)
                sink.push((int)((stopped ? stop_r : carry) + 1u));
        }
    }
}

// ---------------------------------------------------------------------------------------------------------
// Ascending bitonic sort of a[0..n) by the whole block; a must have room for next_pow2(n) ints.
__device__ __forceinline__ int next_pow2(int n) { int p = 1; while (p < n) p <<= 1; return p; }

__device__ inline void block_bitonic_sort(int *a, int n) {
    const int P = next_pow2(n);
    for (int i = n + threadIdx.x; i < P; i += blockDim.x) a[i] = 0x7fffffff;
    __syncthreads();
    for (int k = 2; k <= P; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int t = threadIdx.x; t < P; t += blockDim.x) {
                const int u = t ^ j;
                if (u > t) {
                    const int x = a[t], y = a[u];
                    const bool asc = (t & k) == 0;
                    if ((x > y) == asc) { a[t] = y; a[u] = x; }
                }
            }
            __syncthreads();
        }
    }
}

// ---------------------------------------------------------------------------------------------------------
// consensus_pos on a sorted segment.  Warp 0 runs the leftward scan, warp 1 the rightward scan; the
// left scan's early return wins, then the right scan's, then the nearer best (ties -> right).
struct ScanResult { int early, value, best, dist; };

__device__ inline ScanResult consensus_scan_warp(const int *a, int n, int pos, Tunables t, bool leftward) {
    const int lane = lane_id();
    int maxc = t.min_count - 1, best = -1, dist = 0x7fffffff;
    int start;
    if (leftward) {
        // lower_bound(a, n, pos+25): index of the last element <= x, clamped to 0 (refinement.c:3-10)
        const int x = pos + SV_MIN_LENGTH / 2;
        int lo = 0, hi = n;                                   // first index with a[idx] > x
        while (lo < hi) { const int m = (lo + hi) >> 1; if (a[m] > x) hi = m; else lo = m + 1; }
        start = lo > 0 ? lo - 1 : 0;
    } else {
        // upper_bound(a, n, pos-25): first element *smaller* than x -- 0 or n-1 on ascending input (refinement.c:12-19)
        start = (a[0] < pos - SV_MIN_LENGTH / 2) ? 0 : n - 1;
    }
    for (int base = 0;; base += 32) {
        const int i = leftward ? start - base - lane : start + base + lane;
        const bool valid = leftward ? (i >= 0) : (i < n);
        const int v = valid ? a[i] : 0;
        const bool inr = valid && iabs_wrap(pos - v) < t.range;
        const unsigned stopm = __ballot_sync(FULL, !inr);
        const int nv = stopm ? (__ffs(stopm) - 1) : 32;       // loop condition fails at the first out-of-range element
        bool pending = lane < nv;
        int count = 0, cand = 0, d = 0;
        if (pending) {
            unsigned long long total = 0;
            if (leftward) {
                int lo = 0, hi = i;                           // first j in [0,i) with a[i] <= a[j] + interval
                while (lo < hi) { const int m = (lo + hi) >> 1; if (v <= a[m] + t.interval) hi = m; else lo = m + 1; }
                count = 1 + (i - lo);
                for (int j = lo; j <= i; ++j) total += (unsigned long long)(long long)a[j];
            } else {
                int lo = i + 1, hi = n;                       // first j in (i,n) with a[j] > a[i] + interval
                while (lo < hi) { const int m = (lo + hi) >> 1; if (a[m] <= v + t.interval) lo = m + 1; else hi = m; }
                count = lo - i;
                for (int j = i; j < lo; ++j) total += (unsigned long long)(long long)a[j];
            }
            cand = (int)((total + (unsigned long long)(long long)(count / 2)) / (unsigned long long)(long long)count);
            d = iabs_wrap(pos - cand);
        }
        // sequential part of the reference loop, resolved lane by lane in scan order
        while (true) {
            const unsigned m = __ballot_sync(FULL, pending && count > maxc);
            if (!m) break;
            const int l = __ffs(m) - 1;
            const int c = __shfl_sync(FULL, count, l), cd = __shfl_sync(FULL, cand, l), dd = __shfl_sync(FULL, d, l);
            if (dd < t.interval) return ScanResult{1, cd, best, dist};
            if (dd < dist) { maxc = c; best = cd; dist = dd; }
            pending = pending && lane > l;
        }
        if (nv < 32) break;
    }
    return ScanResult{0, -1, best, dist};
}

// consensus_pos for n <= 32 sorted values held one per lane (lane i: a[i], i < n): both scans of refinement.c:56-98 without
// shared memory.  The cluster behind an anchor is a contiguous run of the sorted values (the j loops stop at the first value
// more than `interval` away), so its bounds come from a 5-step search over the lanes and its sum from one 64-bit prefix scan;
// the reference's sequential max_count / distance bookkeeping is then replayed in scan order over the lanes that qualify.
__device__ __forceinline__ int consensus_pos_lanes(int v, int n, int pos, Tunables t) {
    const int lane = lane_id();
    const bool valid = lane < n;
    // inclusive prefix sums of the values as the reference accumulates them: uint64_t total += int (sign-extended)
    unsigned long long P = valid ? (unsigned long long)(long long)v : 0ull;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const unsigned long long o = __shfl_up_sync(FULL, P, d); if (lane >= d) P += o; }
    const unsigned inr = __ballot_sync(FULL, valid && iabs_wrap(pos - v) < t.range);
    // cluster bounds: lo = first j <= lane with a[lane] <= a[j] + interval ; hi = last j >= lane with a[j] <= a[lane] + interval
    int lo = 0, lo_hi = lane, hi = lane, hi_hi = n - 1;           // search ranges [lo, lo_hi] and [hi, hi_hi] (answers inside)
#pragma unroll
    for (int k = 0; k < 5; ++k) {
        const int m1 = (lo + lo_hi) >> 1, m2 = (hi + hi_hi + 1) >> 1;
        const int a1 = __shfl_sync(FULL, v, m1 & 31), a2 = __shfl_sync(FULL, v, m2 & 31);
        if (lo < lo_hi) { if (v <= a1 + t.interval) lo_hi = m1; else lo = m1 + 1; }
        if (hi < hi_hi) { if (a2 <= v + t.interval) hi = m2; else hi_hi = m2 - 1; }
    }
    const int loc = lo_hi, hic = valid ? hi : lane;               // converged (5 steps cover 32 lanes)
    const unsigned long long Plo = __shfl_sync(FULL, P, (loc - 1) & 31), Phi = __shfl_sync(FULL, P, hic & 31), Pm1 = __shfl_up_sync(FULL, P, 1);
    int best_l = -1, dist_l = 0x7fffffff, best_r = -1, dist_r = 0x7fffffff;
    {   // ---- leftward scan from lower_bound(pos + 25): index of the last value <= x, clamped to 0 (refinement.c:3-10,56)
        const unsigned gt = __ballot_sync(FULL, valid && v > pos + SV_MIN_LENGTH / 2);
        const int first = gt ? __ffs(gt) - 1 : n;
        const int start = gt ? (first > 0 ? first - 1 : 0) : n - 1;
        const unsigned upto = start >= 31 ? 0xffffffffu : ((2u << start) - 1u);
        const unsigned stop = ~inr & upto;                        // the loop ends at the first out-of-range value going down
        const unsigned run = upto & (stop ? ~((2u << (31 - __clz(stop))) - 1u) : 0xffffffffu);
        const int count = lane - loc + 1;
        const unsigned long long total = P - (loc > 0 ? Plo : 0ull);
        const int cand = (int)((total + (unsigned long long)(long long)(count / 2)) / (unsigned long long)(long long)count);
        const int d = iabs_wrap(pos - cand);
        int maxc = t.min_count - 1;
        unsigned pending = run;
        while (true) {
            const unsigned m = __ballot_sync(FULL, ((pending >> lane) & 1u) && count > maxc);
            if (!m) break;
            const int l = 31 - __clz(m);                          // scan order: highest lane first
            const int c = __shfl_sync(FULL, count, l), cd = __shfl_sync(FULL, cand, l), dd = __shfl_sync(FULL, d, l);
            if (dd < t.interval) return cd;                       // refinement.c:67-68
            if (dd < dist_l) { maxc = c; best_l = cd; dist_l = dd; }
            pending &= (1u << l) - 1u;
        }
    }
    {   // ---- rightward scan from upper_bound(pos - 25): 0 or n - 1 on ascending input (refinement.c:12-19,78)
        const int a0 = __shfl_sync(FULL, v, 0);
        const int start = (a0 < pos - SV_MIN_LENGTH / 2) ? 0 : n - 1;
        const unsigned from = start >= 32 ? 0u : ~((1u << start) - 1u);
        const unsigned stop = ~inr & from;                        // first out-of-range (or invalid) lane going up
        const unsigned run = from & (stop ? ((1u << (__ffs(stop) - 1)) - 1u) : 0xffffffffu);
        const int count = hic - lane + 1;
        const unsigned long long total = Phi - (lane > 0 ? Pm1 : 0ull);
        const int cand = (int)((total + (unsigned long long)(long long)(count / 2)) / (unsigned long long)(long long)count);
        const int d = iabs_wrap(pos - cand);
        int maxc = t.min_count - 1;
        unsigned pending = run;
        while (true) {
            const unsigned m = __ballot_sync(FULL, ((pending >> lane) & 1u) && count > maxc);
            if (!m) break;
            const int l = __ffs(m) - 1;                           // scan order: lowest lane first
            const int c = __shfl_sync(FULL, count, l), cd = __shfl_sync(FULL, cand, l), dd = __shfl_sync(FULL, d, l);
            if (dd < t.interval) return cd;                       // refinement.c:89-90
            if (dd < dist_r) { maxc = c; best_r = cd; dist_r = dd; }
            pending &= ~((2u << l) - 1u);
        }
    }
    return dist_l < dist_r ? best_l : best_r;                     // refinement.c:100
}

// a: sorted, n >= 1, n >= min_count.  s_res: 2 ScanResult slots in shared memory.  Needs >= 2 warps.
__device__ inline int block_consensus(const int *a, int n, int pos, Tunables t, ScanResult *s_res) {
    const int warp = threadIdx.x >> 5;
    if (warp < 2) {
        ScanResult r = consensus_scan_warp(a, n, pos, t, warp == 0);
        if (lane_id() == 0) s_res[warp] = r;
    }
    __syncthreads();
    const ScanResult L = s_res[0], R = s_res[1];
    __syncthreads();
    if (L.early) return L.value;
    if (R.early) return R.value;
    return L.dist < R.dist ? L.best : R.best;
}

// ---------------------------------------------------------------------------------------------------------
// sliding_window_ins per-tile scan on a sorted segment (sliding_window.c:70-84): for i = 0, slide, 2*slide, ...
// support_i = #{e >= i : a[e] - a[i] <= windowSize (contiguous)}; the FIRST i with the largest support that is
// >= min_count wins; candidate = (int32 wrapping sum + support/2) / support.  Returns support (0 = none).
__device__ inline int block_slide(const int *a, int n, int window_size, int slide, int min_count, int *s_red, int *best_out) {
    // packed key: support in the high bits, inverted anchor index in the low bits -> max() picks the first best anchor
    unsigned long long key = 0;
    for (int i = threadIdx.x * slide; i < n && i >= 0; i += blockDim.x * slide) {
        const int v = a[i];
        int lo = i, hi = n;                                   // first e with a[e] - a[i] > windowSize
        while (lo < hi) { const int m = (lo + hi) >> 1; if (a[m] - v <= window_size) lo = m + 1; else hi = m; }
        const int support = lo - i;
        const unsigned long long k = ((unsigned long long)(unsigned)support << 32) | (unsigned)(0x7fffffff - i);
        if (k > key) key = k;
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        const unsigned long long o = __shfl_xor_sync(FULL, key, d);
        if (o > key) key = o;
    }
    unsigned long long *s_key = reinterpret_cast<unsigned long long *>(s_red);
    if (lane_id() == 0) s_key[threadIdx.x >> 5] = key;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long k = 0;
        for (int w = 0; w < (int)((blockDim.x + 31) >> 5); ++w) if (s_key[w] > k) k = s_key[w];
        s_key[0] = k;
    }
    __syncthreads();
    key = s_key[0];
    __syncthreads();
    const int support = (int)(key >> 32);
    if (support < min_count || support <= 0) return 0;
    const int i0 = 0x7fffffff - (int)(unsigned)(key & 0xffffffffu);
    int part = 0;
    for (int j = i0 + threadIdx.x; j < i0 + support; j += blockDim.x) part += a[j];
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) part += __shfl_xor_sync(FULL, part, d);
    if (lane_id() == 0) s_red[threadIdx.x >> 5] = part;
    __syncthreads();
    int sum = 0;
    for (int w = 0; w < (int)((blockDim.x + 31) >> 5); ++w) sum += s_red[w];
    __syncthreads();
    *best_out = (sum + support / 2) / support;
    return support;
}

}  // namespace svtk
