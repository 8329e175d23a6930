// svtk_kernels.cu -- the sm_100a kernels of the audt refinement path and their launchers.
//
//   k_max_span          one pass over read metadata: max(end - pos), the planner's look-back bound
//   k_plan_windows      per window: which reads the reference's BAM iterator would yield, as an index range
//                       (sam_itr_queryi/sam_itr_next semantics, hts.c:3049-3090, 3905-3921)
//   k_refine_fused      one CTA per window: CIGAR walk (one warp per read, uint4 loads) -> signatures in shared
//                       memory -> bitonic sort -> consensus_pos.  Signatures never touch HBM.
//   k_refine_overflow   windows whose signature count exceeded the fast path's shared-memory segment: same
//                       pipeline with a 128 KB segment, or a global-memory segment beyond that
//   k_consensus_segments / k_slide_tiles / k_window_signatures   parity aids + sliding_window_ins
//
// HBM-bound integer work: grids are one CTA per window (thousands of windows per batch, >> 148 SMs), loads are
// 128-bit and coalesced, every byte of CIGAR is read once per (window, read) pair like the reference does.
#include "svtk_device.cuh"
#include "svtk_launch.h"

namespace svtk {

// ---------------------------------------------------------------------------------------------------------
__global__ void k_max_span(const uint32_t *__restrict__ pos, const uint32_t *__restrict__ end, long long n, uint32_t *out) {
    uint32_t m = 1;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const uint32_t s = end[i] - pos[i];
        m = s > m ? s : m;
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) { const uint32_t o = __shfl_xor_sync(FULL, m, d); m = o > m ? o : m; }
    if (lane_id() == 0) atomicMax(out, m);
}

// first index in [lo,hi) with pos[idx] >= key
__device__ __forceinline__ long long lower_bound_pos(const uint32_t *__restrict__ pos, long long lo, long long hi, long long key) {
    while (lo < hi) {
        const long long m = (lo + hi) >> 1;
        if ((long long)__ldg(pos + m) >= key) hi = m; else lo = m + 1;
    }
    return lo;
}

// range.x = first candidate read, range.y = one past the last; candidates still need end > beg (checked by the walker)
__global__ void k_plan_windows(DevReads R, DevWindows W, int2 *__restrict__ range) {
    const int w = blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= W.n_win) return;
    int2 r = make_int2(0, 0);
    const int tid = W.chrom[w] - 1;                          // refinement.c:114 (chrom-1)
    const long long beg = (long long)(uint32_t)(W.start[w] - 1u);   // uint32 arithmetic, then widened
    const long long endq = (long long)(uint32_t)(W.end[w] - 1u);
    if (tid >= 0 && tid < R.n_tid && endq >= beg && W.kind[w] != KIND_POINT) {
        const long long t0 = R.tid_off[tid], t1 = R.tid_off[tid + 1];
        const long long hi = lower_bound_pos(R.pos, t0, t1, endq);                       // yielded reads have pos < endq
        const long long lo = lower_bound_pos(R.pos, t0, hi, beg - (long long)R.max_span + 1);   // and pos + span > beg
        r = make_int2((int)lo, (int)hi);
    }
    range[w] = r;
}

// ---------------------------------------------------------------------------------------------------------
// Block-wide signature collection for one window: tiles of NT reads are filtered (end > beg, n_cigar > 0),
// compacted into shared memory, then consumed one read per warp.
template <int NT>
struct CollectShared {
    uint32_t l_pos[NT];
    uint32_t l_n[NT];
    unsigned long long l_off[NT];
    int n_list;
};

template <int NT, bool STATS>
__device__ __forceinline__ void collect_window(const DevReads &R, int kind, uint32_t w_start, uint32_t w_end, int lo, int hi,
                                               const SigSink &sink, CollectShared<NT> &cs, unsigned long long &pairs,
                                               unsigned long long &words) {
    const uint32_t beg = w_start - 1u;
    const int warp = threadIdx.x >> 5, lane = lane_id();
    constexpr int NW = NT / 32;
    for (int tile = lo; tile < hi; tile += NT) {
        if (threadIdx.x == 0) cs.n_list = 0;
        __syncthreads();
        const int r = tile + (int)threadIdx.x;
        bool ok = false;
        uint32_t rp = 0, rn = 0;
        unsigned long long ro = 0;
        if (r < hi) {
            rn = __ldg(R.n_cigar + r);
            ok = (__ldg(R.end + r) > beg) && rn > 0;
            if (ok) { rp = __ldg(R.pos + r); ro = __ldg(R.cigar_off + r); }
        }
        const unsigned m = __ballot_sync(FULL, ok);
        int wbase = 0;
        if (lane == 0 && m) wbase = atomicAdd(&cs.n_list, __popc(m));
        wbase = __shfl_sync(FULL, wbase, 0);
        if (ok) {
            const int slot = wbase + __popc(m & ((1u << lane) - 1u));
            cs.l_pos[slot] = rp; cs.l_n[slot] = rn; cs.l_off[slot] = ro;
        }
        __syncthreads();
        const int nl = cs.n_list;
        for (int k = warp; k < nl; k += NW)
            walk_read_warp<STATS>(R.cigar, kind, w_start, w_end, cs.l_pos[k], cs.l_n[k], cs.l_off[k], sink, words);
        if (STATS && warp == 0 && lane == 0) pairs += (unsigned long long)nl;
        __syncthreads();
    }
}

// ---------------------------------------------------------------------------------------------------------
// Fast path: one CTA per window, signatures in a CAP-int shared segment.
template <int NT, int CAP, bool STATS>
__global__ void __launch_bounds__(NT) k_refine_fused(DevReads R, DevWindows W, const int2 *__restrict__ range, Tunables t,
                                                     int32_t *__restrict__ result, OverflowList ov, WorkCounters *wc) {
    __shared__ int s_sig[CAP];
    __shared__ int s_count;
    __shared__ CollectShared<NT> cs;
    __shared__ ScanResult s_res[2];
    const int w = blockIdx.x;
    const int2 rg = range[w];
    if (rg.y <= rg.x) {                     // no iterator / no reads / refine_point: consensus_pos(empty) == -1
        if (threadIdx.x == 0) result[w] = -1;
        return;
    }
    if (threadIdx.x == 0) s_count = 0;
    __syncthreads();
    unsigned long long pairs = 0, words = 0;
    const SigSink sink{s_sig, &s_count, CAP};
    collect_window<NT, STATS>(R, W.kind[w], W.start[w], W.end[w], rg.x, rg.y, sink, cs, pairs, words);
    const int n = s_count;
    if (STATS) {
        if (lane_id() == 0 && words) atomicAdd(&wc->words, words);
        if (threadIdx.x == 0) { atomicAdd(&wc->pairs, pairs); atomicAdd(&wc->signatures, (unsigned long long)n); }
    }
    if (n < t.min_count || n <= 0) {        // refinement.c:43
        if (threadIdx.x == 0) result[w] = -1;
        return;
    }
    if (n > CAP) {                          // leave the fast path: remember (window, count) for k_refine_overflow
        if (threadIdx.x == 0) {
            const int slot = atomicAdd(ov.count, 1);
            if (slot < ov.capacity) ov.items[slot] = make_int2(w, n);
            result[w] = -1;
            if (STATS) atomicAdd(&wc->overflow, 1ull);
        }
        return;
    }
    block_bitonic_sort(s_sig, n);
    const int res = block_consensus(s_sig, n, (int)W.pos[w], t, s_res);
    if (threadIdx.x == 0) result[w] = res;
}

// ---------------------------------------------------------------------------------------------------------
// Overflow path: persistent CTAs walk the overflow list.  Segment = dynamic shared memory (cap_smem ints) when
// the window fits, else a slice of the global arena (bump-allocated, next_pow2(n) ints).
template <int NT>
__global__ void __launch_bounds__(NT) k_refine_overflow(DevReads R, DevWindows W, const int2 *__restrict__ range, Tunables t,
                                                        int32_t *__restrict__ result, OverflowList ov, int cap_smem,
                                                        int *arena, unsigned long long arena_ints, unsigned long long *arena_top,
                                                        int *fail_flag) {
    extern __shared__ int s_dyn[];
    __shared__ int s_count;
    __shared__ unsigned long long s_base;
    __shared__ CollectShared<NT> cs;
    __shared__ ScanResult s_res[2];
    const int n_items = min(*ov.count, ov.capacity);
    for (int it = blockIdx.x; it < n_items; it += gridDim.x) {
        const int2 item = ov.items[it];
        const int w = item.x, n_expect = item.y;
        int *seg = s_dyn;
        int cap = cap_smem;
        if (n_expect > cap_smem) {
            if (threadIdx.x == 0) s_base = atomicAdd(arena_top, (unsigned long long)next_pow2(n_expect));
            __syncthreads();
            if (s_base + (unsigned long long)next_pow2(n_expect) > arena_ints) {   // arena exhausted: host retries with a bigger one
                if (threadIdx.x == 0) atomicExch(fail_flag, 1);
                __syncthreads();
                continue;
            }
            seg = arena + s_base;
            cap = n_expect;
        }
        if (threadIdx.x == 0) s_count = 0;
        __syncthreads();
        unsigned long long pairs = 0, words = 0;
        const int2 rg = range[w];
        const SigSink sink{seg, &s_count, cap};
        collect_window<NT, false>(R, W.kind[w], W.start[w], W.end[w], rg.x, rg.y, sink, cs, pairs, words);
        __threadfence_block();
        const int n = s_count;               // == n_expect
        block_bitonic_sort(seg, n);
        const int res = block_consensus(seg, n, (int)W.pos[w], t, s_res);
        if (threadIdx.x == 0) result[w] = res;
        __syncthreads();
    }
}

// ---------------------------------------------------------------------------------------------------------
// consensus_pos on caller-supplied arrays (parity aid for refinement.c:41): one CTA per segment.
template <int NT>
__global__ void __launch_bounds__(NT) k_consensus_segments(const int32_t *__restrict__ values, const long long *__restrict__ offsets,
                                                           const int32_t *__restrict__ pos, Tunables t, int *scratch,
                                                           const long long *__restrict__ scratch_off, int cap_smem,
                                                           int32_t *__restrict__ result) {
    extern __shared__ int s_dyn[];
    __shared__ ScanResult s_res[2];
    const int s = blockIdx.x;
    const long long b = offsets[s];
    const int n = (int)(offsets[s + 1] - b);
    if (n < t.min_count || n <= 0) { if (threadIdx.x == 0) result[s] = -1; return; }
    int *seg = (n <= cap_smem) ? s_dyn : scratch + scratch_off[s];
    for (int i = threadIdx.x; i < n; i += NT) seg[i] = values[b + i];
    __syncthreads();
    block_bitonic_sort(seg, n);
    const int res = block_consensus(seg, n, pos[s], t, s_res);
    if (threadIdx.x == 0) result[s] = res;
}

// ---------------------------------------------------------------------------------------------------------
// sliding_window_ins tiles (sliding_window.c:12-94): one CTA per tile; walk with the INS rule, sort, support scan.
// out[tile] = (best candidate, support), support == 0 when the tile prints nothing.  Tiles whose signatures do
// not fit the segment set *fail_flag (the host retries with a global segment).
template <int NT>
__global__ void __launch_bounds__(NT) k_slide_tiles(DevReads R, DevWindows W, const int2 *__restrict__ range, int window_size,
                                                    int slide, int min_count, int cap, int *gseg, long long gseg_stride,
                                                    int2 *__restrict__ out, int *need) {
    extern __shared__ int s_dyn[];
    __shared__ int s_count;
    __shared__ int s_red[64];
    __shared__ CollectShared<NT> cs;
    const int w = blockIdx.x;
    const int2 rg = range[w];
    int *seg = gseg ? gseg + (long long)w * gseg_stride : s_dyn;
    if (threadIdx.x == 0) s_count = 0;
    __syncthreads();
    unsigned long long pairs = 0, words = 0;
    const SigSink sink{seg, &s_count, cap};
    if (rg.y > rg.x) collect_window<NT, false>(R, KIND_INS, W.start[w], W.end[w], rg.x, rg.y, sink, cs, pairs, words);
    __syncthreads();
    const int n = s_count;
    if (n > cap) { if (threadIdx.x == 0) { atomicMax(need, n); out[w] = make_int2(-1, -1); } return; }
    if (n == 0) { if (threadIdx.x == 0) out[w] = make_int2(-1, 0); return; }
    block_bitonic_sort(seg, n);
    int best = -1;
    const int support = block_slide(seg, n, window_size, slide, min_count, s_red, &best);
    if (threadIdx.x == 0) out[w] = make_int2(support ? best : -1, support);
}

// ---------------------------------------------------------------------------------------------------------
// Raw signature multiset of ONE window into a global buffer, sorted ascending when it fits (parity aid).
template <int NT>
__global__ void __launch_bounds__(NT) k_window_signatures(DevReads R, DevWindows W, const int2 *__restrict__ range, int *buf,
                                                          int cap, int *n_out) {
    __shared__ int s_count;
    __shared__ CollectShared<NT> cs;
    const int2 rg = range[0];
    if (threadIdx.x == 0) s_count = 0;
    __syncthreads();
    unsigned long long pairs = 0, words = 0;
    const SigSink sink{buf, &s_count, cap};
    if (rg.y > rg.x) collect_window<NT, false>(R, W.kind[0], W.start[0], W.end[0], rg.x, rg.y, sink, cs, pairs, words);
    __syncthreads();
    __threadfence_block();
    const int n = s_count;
    if (threadIdx.x == 0) *n_out = n;
    if (n > 0 && next_pow2(n) <= cap) block_bitonic_sort(buf, n);
}

// =========================================================================================================
// launchers (called from svtk_capi.cpp through svtk_launch.h; no CUDA types leak beyond cudaStream_t)
constexpr int FUSED_NT = 128;
constexpr int FUSED_CAP = 1024;
constexpr int BIG_NT = 1024;

static DevReads to_dev(const LaunchReads &r) {
    DevReads d;
    d.n_tid = r.n_tid; d.n_reads = r.n_reads; d.tid_off = r.tid_off; d.pos = r.pos; d.end = r.end; d.n_cigar = r.n_cigar;
    d.cigar_off = r.cigar_off; d.cigar = r.cigar; d.max_span = r.max_span; d.tiles = r.tiles; d.tile_off = r.tile_off;
    return d;
}
static DevWindows to_dev(const LaunchWindows &w) {
    DevWindows d;
    d.n_win = w.n_win; d.chrom = w.chrom; d.start = w.start; d.end = w.end; d.pos = w.pos; d.kind = w.kind;
    return d;
}

int big_segment_ints() {
    // dynamic shared memory of the overflow kernel next to its static arrays (227 KB opt-in per CTA); a power of two
    // because the bitonic sort pads a segment to next_pow2(n)
    return 32768;
}

cudaError_t launch_max_span(const uint32_t *pos, const uint32_t *end, long long n, uint32_t *out, cudaStream_t s) {
    if (n <= 0) return cudaSuccess;
    const int blocks = (int)((n + 1023) / 1024 < 148 * 8 ? (n + 1023) / 1024 : 148 * 8);
    k_max_span<<<blocks, 256, 0, s>>>(pos, end, n, out);
    return cudaGetLastError();
}

cudaError_t launch_plan(const LaunchReads &r, const LaunchWindows &w, int2 *range, cudaStream_t s) {
    if (w.n_win <= 0) return cudaSuccess;
    k_plan_windows<<<(w.n_win + 127) / 128, 128, 0, s>>>(to_dev(r), to_dev(w), range);
    return cudaGetLastError();
}

cudaError_t launch_refine_fused(const LaunchReads &r, const LaunchWindows &w, const int2 *range, int range_, int interval,
                                int min_count, int32_t *result, int *ov_count, int2 *ov_items, int ov_capacity,
                                unsigned long long *counters, cudaStream_t s) {
    if (w.n_win <= 0) return cudaSuccess;
    const Tunables t{range_, interval, min_count};
    const OverflowList ov{ov_count, ov_items, ov_capacity};
    if (counters)
        k_refine_fused<FUSED_NT, FUSED_CAP, true><<<w.n_win, FUSED_NT, 0, s>>>(to_dev(r), to_dev(w), range, t, result, ov,
                                                                                reinterpret_cast<WorkCounters *>(counters));
    else
        k_refine_fused<FUSED_NT, FUSED_CAP, false><<<w.n_win, FUSED_NT, 0, s>>>(to_dev(r), to_dev(w), range, t, result, ov, nullptr);
    return cudaGetLastError();
}

cudaError_t launch_refine_overflow(const LaunchReads &r, const LaunchWindows &w, const int2 *range, int range_, int interval,
                                   int min_count, int32_t *result, int *ov_count, int2 *ov_items, int ov_capacity, int n_items,
                                   int *arena, unsigned long long arena_ints, unsigned long long *arena_top, int *fail_flag,
                                   cudaStream_t s) {
    static unsigned long long attr_seen = 0;
    const int cap = big_segment_ints();
    if (first_use_on_device(attr_seen)) {
        cudaError_t e = cudaFuncSetAttribute(k_refine_overflow<BIG_NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, cap * 4);
        if (e != cudaSuccess) return e;
    }
    const Tunables t{range_, interval, min_count};
    const OverflowList ov{ov_count, ov_items, ov_capacity};
    const int grid = n_items < 148 ? n_items : 148;
    k_refine_overflow<BIG_NT><<<grid, BIG_NT, cap * 4, s>>>(to_dev(r), to_dev(w), range, t, result, ov, cap, arena, arena_ints,
                                                           arena_top, fail_flag);
    return cudaGetLastError();
}

cudaError_t launch_consensus_segments(const int32_t *values, const long long *offsets, const int32_t *pos, int n_seg, int range_,
                                      int interval, int min_count, int *scratch, const long long *scratch_off, int32_t *result,
                                      cudaStream_t s) {
    if (n_seg <= 0) return cudaSuccess;
    constexpr int NT = 256;
    const int cap = 8192;
    const Tunables t{range_, interval, min_count};
    k_consensus_segments<NT><<<n_seg, NT, cap * 4, s>>>(values, offsets, pos, t, scratch, scratch_off, cap, result);
    return cudaGetLastError();
}
int consensus_segment_smem_ints() { return 8192; }

cudaError_t launch_slide_tiles(const LaunchReads &r, const LaunchWindows &w, const int2 *range, int window_size, int slide,
                               int min_count, int *gseg, long long gseg_stride, int2 *out, int *need, cudaStream_t s) {
    if (w.n_win <= 0) return cudaSuccess;
    constexpr int NT = 256;
    const int cap_smem = 8192;
    if (gseg)
        k_slide_tiles<NT><<<w.n_win, NT, 0, s>>>(to_dev(r), to_dev(w), range, window_size, slide, min_count, (int)gseg_stride, gseg,
                                                 gseg_stride, out, need);
    else
        k_slide_tiles<NT><<<w.n_win, NT, cap_smem * 4, s>>>(to_dev(r), to_dev(w), range, window_size, slide, min_count, cap_smem,
                                                            nullptr, 0, out, need);
    return cudaGetLastError();
}

cudaError_t launch_window_signatures(const LaunchReads &r, const LaunchWindows &w, const int2 *range, int *buf, int cap, int *n_out,
                                     cudaStream_t s) {
    k_window_signatures<BIG_NT><<<1, BIG_NT, 0, s>>>(to_dev(r), to_dev(w), range, buf, cap, n_out);
    return cudaGetLastError();
}

}  // namespace svtk
