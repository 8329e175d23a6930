// svtk_poa.cu -- K3 on the device: k_poa_sets (one warp-sized CTA per read set, persistent workers pulling
// sets from an atomic queue) and the host driver behind svtk_poa_consensus (include/svtk.h).
//
// Work unit = one INS record's read set (abpoa_msa, abPOA/src/abpoa_align.c:401).  Sets are independent, reads
// inside a set are strictly sequential (the graph grows), so the device is filled with sets, not with cells of
// one alignment: up to 32 resident CTAs per SM x 148 SMs read sets are in flight, each warp alternating between
// its column-parallel row DP and its lane-0 graph phases (svtk_poa_core.cuh) while the other warps of the SM
// hide the latency.  Every worker owns one scratch block in HBM (graph SoA, edge pool, band-only DP rows with
// all five score planes for the backtrack, cigar); blocks are sized per size class (read sets binned by
// ceil(log2(longest read))) so that a 100 bp set does not reserve a 10 kb set's 200 MB.  A set that outgrows
// its estimate reports which resource ran out and is rerun with a larger block -- never silently truncated.
#include "svtk_ctx.h"
#include "svtk_poa_core.cuh"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

namespace svtk {

using namespace svtk_poa;

struct PoaCaps {
    int node_cap, edge_cap, cigar_cap;
    long long cell_bytes;
};

// layout of one worker's scratch block (byte offsets, 16-byte aligned sections)
struct PoaLayout {
    long long base, ints, row_off, edge_id, edge_w, edge_idx, rowmeta, cigar, cells, total;
};
constexpr int POA_INT_ARRAYS = 17 + ALN_MAX;

__host__ __device__ inline long long a16(long long x) { return (x + 15) & ~15ll; }
__host__ __device__ inline PoaLayout poa_layout(const PoaCaps &c) {
    PoaLayout L;
    long long o = 0;
    L.base = o; o = a16(o + c.node_cap);
    L.ints = o; o = a16(o + 4ll * POA_INT_ARRAYS * c.node_cap);
    L.row_off = o; o = a16(o + 8ll * c.node_cap);
    L.edge_id = o; o = a16(o + 4ll * c.edge_cap);
    L.edge_w = o; o = a16(o + 4ll * c.edge_cap);
    L.edge_idx = o; o = a16(o + 4ll * c.edge_cap);
    L.rowmeta = o; o = a16(o + 48ll * c.node_cap);           // rowmeta, rowout, rowinfo
    L.cigar = o; o = a16(o + 8ll * c.cigar_cap);
    L.cells = o; o = a16(o + c.cell_bytes);
    L.total = (o + 255) & ~255ll;
    return L;
}

// One size class of a launch: its slice of the item list, its cursor, its workers and their scratch blocks.
struct PoaClass {
    int item_begin, item_end;          // slice of PoaLaunch::items (heaviest first)
    int worker_begin, worker_end;      // CTAs that own a block of this class
    long long scratch_off;             // byte offset of this class's first block
    PoaCaps caps;                      // >= the caps of every smaller class: a worker can also take their sets
};

struct PoaLaunch {
    int n_classes;
    const PoaClass *classes;           // largest reads first
    const int *items;                  // read-set ids
    int *cursors;                      // one atomic cursor per class
    const long long *set_off, *seq_off;
    const uint8_t *bases;
    uint8_t *cons;
    int32_t *cov;                      // may be null
    const long long *cons_off;
    int32_t *cons_len, *status;
    unsigned long long *cells;         // [0] band cells, [1] full-matrix equivalent, [2..5] SM cycles per phase
    unsigned char *scratch;
    PoaTrace trace;
    int have_trace;
    unsigned long long *times;         // SVTK_POA_DEBUG: per read set {start, end} of %globaltimer (ns), then {peak cells, nodes | edges << 32}; else null
    int n_sets_dbg;
};

// NW = 1: one warp per read set (the throughput shape).  NW = 8: the whole 256-thread CTA shares each DP row and the
// parallel parts of the sort -- used for the few heaviest sets, whose single-warp latency would be the batch's tail.
#ifndef SVTK_POA_MINB
#define SVTK_POA_MINB 16
#endif
template <int NW>
__global__ void __launch_bounds__(32 * NW, NW == 1 ? SVTK_POA_MINB : 1) k_poa_sets(PoaLaunch L) {
    __shared__ PoaState s;
    __shared__ PoaXchg xch;
    __shared__ int s_item, s_cls;
    const int lane = threadIdx.x;
    if (lane == 0) {
        int k = 0;
        while (k + 1 < L.n_classes && (int)blockIdx.x >= L.classes[k].worker_end) ++k;
        s_cls = k;
        const PoaClass cl = L.classes[k];
        const PoaLayout lay = poa_layout(cl.caps);
        unsigned char *blk = L.scratch + cl.scratch_off + (long long)((int)blockIdx.x - cl.worker_begin) * lay.total;
        const int N = cl.caps.node_cap;
        s.node_cap = N; s.edge_cap = cl.caps.edge_cap; s.cigar_cap = cl.caps.cigar_cap; s.cell_bytes = cl.caps.cell_bytes;
        s.base = blk + lay.base;
        int *p = reinterpret_cast<int *>(blk + lay.ints);
        s.in_off = p; p += N; s.in_n = p; p += N; s.in_cap = p; p += N; s.out_off = p; p += N; s.out_n = p; p += N; s.out_cap = p; p += N;
        s.aln_n = p; p += N; s.aln_id = p; p += (long long)ALN_MAX * N; s.n_read = p; p += N;
        s.idx2id = p; p += N; s.id2idx = p; p += N; s.max_left = p; p += N; s.max_right = p; p += N; s.remain = p; p += N;
        s.deg = p; p += N; s.queue = p; p += N; s.dp_beg = p; p += N; s.dp_end = p; p += N;
        s.row_off = reinterpret_cast<long long *>(blk + lay.row_off);
        s.edge_id = reinterpret_cast<int *>(blk + lay.edge_id);
        s.edge_w = reinterpret_cast<int *>(blk + lay.edge_w);
        s.edge_idx = reinterpret_cast<int *>(blk + lay.edge_idx);
        s.rowmeta = reinterpret_cast<int4 *>(blk + lay.rowmeta);
        s.rowout = s.rowmeta + N;
        s.rowinfo = s.rowmeta + 2 * (long long)N;
        s.cigar = reinterpret_cast<int2 *>(blk + lay.cigar);
        s.cells = blk + lay.cells;
    }
    __syncthreads();
    // own class first, then every smaller one (their sets fit this worker's block): no CTA idles while sets remain
    for (int k = s_cls; k < L.n_classes; ++k) {
        const int i0 = L.classes[k].item_begin, n_items = L.classes[k].item_end - i0;
        while (true) {
            if (lane == 0) s_item = atomicAdd(L.cursors + k, 1);
            __syncthreads();
            const int it = s_item;
            __syncthreads();
            if (it >= n_items) break;
            const int sv = L.items[i0 + it];
            const long long co = L.cons_off[sv];
            const int cap = (int)(L.cons_off[sv + 1] - co);
            if (L.times && lane == 0) { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); L.times[2 * sv] = t; }
            const int len = poa_run_set<NW>(s, xch, L.set_off[sv], L.set_off[sv + 1], L.seq_off, L.bases, L.cons + co, L.cov ? L.cov + co : nullptr, cap,
                                        L.have_trace ? &L.trace : nullptr);
            if (lane == 0) {
                if (L.times) { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); L.times[2 * sv + 1] = t;
                               L.times[2 * L.n_sets_dbg + 2 * sv] = (unsigned long long)s.cells_peak; L.times[2 * L.n_sets_dbg + 2 * sv + 1] = (unsigned long long)s.node_n | (unsigned long long)s.edge_top << 32; }
                L.status[sv] = s.status;
                L.cons_len[sv] = len;
                if (!s.status) {
                    atomicAdd(&L.cells[0], (unsigned long long)s.cells_done); atomicAdd(&L.cells[1], (unsigned long long)s.cells_full);
                    for (int k = 0; k < 4; ++k) atomicAdd(&L.cells[2 + k], (unsigned long long)s.t_phase[k]);
                }
            }
            __syncthreads();
        }
    }
}

#define CU(call)                                                                                              \
    do {                                                                                                      \
        cudaError_t e_ = (call);                                                                              \
        if (e_ != cudaSuccess) return fail(SVTK_ERR_CUDA, "%s: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

struct SetInfo {
    int id, n, maxlen, minlen, cls;
    long long total;
    double work;
};

static PoaCaps caps_for(const SetInfo &si, double scale) {
    PoaCaps c;
    const long long worst_nodes = si.total + 2;
    long long nodes = (long long)(scale * 5.0 * si.maxlen) + 256;
    if (nodes > worst_nodes) nodes = worst_nodes;
    if (nodes < 8) nodes = 8;
    c.node_cap = (int)nodes;
    c.edge_cap = (int)std::min<long long>(8 * nodes + 4ll * si.n + 1024, 0x7fffffffll);
    c.cigar_cap = (int)std::min<long long>(nodes + si.maxlen + 16, 0x7fffffffll);
    const int w = EXTRA_B + (int)(EXTRA_F * (float)si.maxlen);
    // band cells of one alignment: rows (graph nodes; ~3.4x the template at 10 % error) x band width
    // (2w + 1 + arg-max spread + how far qlen - remain drifts from the diagonal, bounded by the length spread)
    long long width = (long long)(scale * (2 * w + 48 + (si.maxlen - si.minlen)));
    if (width > si.maxlen + 1) width = si.maxlen + 1;
    const long long dp_rows = std::min<long long>(nodes, (long long)(scale * 4.0 * si.maxlen) + 256);
    const long long len = std::max<long long>(si.maxlen, nodes);
    const int sc_bytes = (std::max<long long>(si.maxlen * MATCH, len * GAP_E1 + GAP_O1) <= 32767 - MISMATCH - GAP_OE1 - GAP_OE2) ? 2 : 4;
    // rows x width is what a set can reach; the C4-like sets (8 % errors, 48 reads) peak at 0.54 of it (profiles/r03m: the most any
    // set of a class used).  3/4 of the product leaves 40 % on top of that; a set that still outgrows its block is rerun with a
    // 4x block (never truncated), and the scratch budget is what limits how many heavy sets run side by side.
    c.cell_bytes = a16((long long)(0.75 * (double)(dp_rows * width)) * NPL * sc_bytes);
    return c;
}

static PoaCaps caps_max(const PoaCaps &a, const PoaCaps &b) {
    PoaCaps c;
    c.node_cap = std::max(a.node_cap, b.node_cap); c.edge_cap = std::max(a.edge_cap, b.edge_cap);
    c.cigar_cap = std::max(a.cigar_cap, b.cigar_cap); c.cell_bytes = std::max(a.cell_bytes, b.cell_bytes);
    return c;
}

// Size classes, workers and scratch of one launch over items[0 .. n) (sorted: class descending, work descending).
// W = CTAs wanted at most.  Fills cls (item ranges relative to item_base, workers and scratch offsets relative to 0).
static int plan_group(const std::vector<SetInfo> &info, const int *items, size_t n, int item_base, double scale, long long W,
                      long long budget, std::vector<PoaClass> &cls, long long &workers, long long &bytes_total) {
    cls.clear(); workers = 0; bytes_total = 0;
    if (n == 0) return 0;
    std::vector<double> cwork;
    for (size_t i0 = 0; i0 < n;) {
        size_t i1 = i0;
        PoaCaps caps = caps_for(info[items[i0]], scale);
        double wk = 0;
        while (i1 < n && info[items[i1]].cls == info[items[i0]].cls) { caps = caps_max(caps, caps_for(info[items[i1]], scale)); wk += info[items[i1]].work; ++i1; }
        PoaClass pc{};
        pc.item_begin = item_base + (int)i0; pc.item_end = item_base + (int)i1; pc.caps = caps;
        cls.push_back(pc); cwork.push_back(wk);
        i0 = i1;
    }
    const int K = (int)cls.size();
    for (int k = K - 2; k >= 0; --k) cls[k].caps = caps_max(cls[k].caps, cls[k + 1].caps);   // stealing downwards must fit
    W = std::min<long long>(W, (long long)n);
    double tot_work = 0;
    for (double x : cwork) tot_work += x;
    // Workers per class.  A set's time is ~ proportional to its work and nothing shortens the heaviest set, so the batch cannot
    // end before T = (heaviest set's work) x (sets the busiest worker runs one after the other).  A class therefore NEEDS
    // ceil(its work / T) workers -- with fewer, its sets queue up behind each other and become the batch's tail (measured: a class
    // of 479 sets of 2-6 s each on 209 workers ended at 10.9 s when the heaviest set of the batch took 5.1 s) -- and gains
    // nothing from more than one worker per set.  Start from T = the heaviest set (x 0.75: estimates are rough, spare
    // workers steal from the smaller classes anyway) and stretch T until the scratch fits the budget; what is left of the
    // budget and of the W CTAs goes to the classes with the largest reads first.
    std::vector<long long> nw(K), bytes(K), nk(K);
    double wmax = 1.0;
    for (size_t i = 0; i < n; ++i) wmax = std::max(wmax, info[items[i]].work);
    long long min_mem = 0;
    for (int k = 0; k < K; ++k) {
        bytes[k] = poa_layout(cls[k].caps).total;
        if (bytes[k] > budget) return fail(SVTK_ERR_ALLOC, "poa: one worker needs %lld bytes of scratch, budget %lld", bytes[k], budget);
        nk[k] = cls[k].item_end - cls[k].item_begin;
        min_mem += bytes[k];
    }
    if (min_mem > budget) return fail(SVTK_ERR_ALLOC, "poa: scratch for one worker per size class (%lld bytes) exceeds the budget %lld", min_mem, budget);
    long long mem = 0, used = 0;
    for (double T = 0.75 * wmax;; T *= 1.15) {
        mem = 0; used = 0;
        for (int k = 0; k < K; ++k) {
            nw[k] = std::max<long long>(1, std::min<long long>(nk[k], (long long)std::ceil(cwork[k] / T)));
            mem += nw[k] * bytes[k]; used += nw[k];
        }
        if (mem <= budget && used <= std::max<long long>(W, K)) break;
        if (T > 4.0 * tot_work + 4.0 * wmax) break;            // every class is down to one worker: min_mem fits (checked above)
    }
    for (int k = 0; k < K && used < W; ++k) {                  // what is left: largest reads first
        const long long room = bytes[k] > 0 ? (budget - mem) / bytes[k] : (nk[k] - nw[k]);
        const long long add = std::max<long long>(0, std::min<long long>(std::min<long long>(nk[k] - nw[k], W - used), room));
        nw[k] += add; used += add; mem += add * bytes[k];
    }
    long long off = 0;
    for (int k = 0; k < K; ++k) {
        cls[k].worker_begin = (int)workers; workers += nw[k]; cls[k].worker_end = (int)workers;
        cls[k].scratch_off = off; off += nw[k] * bytes[k];
    }
    bytes_total = off;
    return 0;
}

}  // namespace svtk

using namespace svtk;

extern "C" int svtk_poa_consensus(svtk_ctx *ctx, const svtk_poa_batch *b, uint8_t *cons, const int64_t *cons_off, int32_t *cons_len,
                                  int32_t *cons_cov, int32_t *set_status, const svtk_poa_trace *trace, svtk_poa_stats *stats) {
    Ctx *c = reinterpret_cast<Ctx *>(ctx);
    if (!c || !b || !cons || !cons_off || !cons_len) return fail(SVTK_ERR_ARG, "null argument");
    if (b->n_sets < 0 || (b->n_sets > 0 && (!b->set_off || !b->seq_off || !b->bases))) return fail(SVTK_ERR_ARG, "bad poa batch");
    if (stats) memset(stats, 0, sizeof(*stats));
    const int ns = b->n_sets;
    if (ns == 0) return 0;
    if (int rc = set_device(c)) return rc;
    const long long n_reads = b->set_off[ns];
    if (b->set_off[0] != 0 || n_reads < 0) return fail(SVTK_ERR_ARG, "poa batch: set_off must start at 0");
    if (b->seq_off[0] != 0) return fail(SVTK_ERR_ARG, "poa batch: seq_off must start at 0");
    const long long n_bases = b->seq_off[n_reads];
    if (trace && !(trace->best_score && trace->node_n && trace->n_ops && trace->ops_hash && trace->cells))
        return fail(SVTK_ERR_ARG, "poa trace: all five arrays are required");

    std::vector<SetInfo> info(ns);
    for (int s = 0; s < ns; ++s) {
        SetInfo &si = info[s];
        si.id = s; si.n = (int)(b->set_off[s + 1] - b->set_off[s]); si.maxlen = 0; si.minlen = 0x7fffffff; si.total = 0;
        if (si.n < 0) return fail(SVTK_ERR_ARG, "poa batch: set_off not monotone at set %d", s);
        if (cons_off[s + 1] < cons_off[s]) return fail(SVTK_ERR_ARG, "cons_off not monotone at set %d", s);
        double work = 0;
        for (long long r = b->set_off[s]; r < b->set_off[s + 1]; ++r) {
            const long long l = b->seq_off[r + 1] - b->seq_off[r];
            if (l <= 0 || l > (1 << 28)) return fail(SVTK_ERR_ARG, "poa batch: read %lld of set %d has length %lld (must be 1 .. 2^28)", r, s, l);
            si.maxlen = std::max(si.maxlen, (int)l); si.minlen = std::min(si.minlen, (int)l); si.total += l;
            work += (double)l;
        }
        if (si.n == 0) { si.minlen = 0; }
        si.work = work * std::max(si.maxlen, 1);
        int cls = 0;
        while ((1 << cls) < std::max(si.maxlen, 1)) ++cls;
        si.cls = cls;                                         // octave of the longest read; refined below
    }
    // Size classes: a worker's scratch block is sized for the largest set of its class, and a block for 8 192-base reads is 4x
    // what a set of 4 100-base reads needs, so octaves waste most of the memory of the heavy classes (and memory is what limits
    // how many heavy sets run side by side).  Eighths of an octave (block within ~1.2x of the need) unless the batch spans so many
    // octaves that the class table would overflow.
    {
        int lo = 64, hi = 0;
        for (int s = 0; s < ns; ++s) if (info[s].n > 0) { lo = std::min(lo, info[s].cls); hi = std::max(hi, info[s].cls); }
        int sub = 8;
        while (sub > 1 && (hi - lo + 1) * sub > 80) sub >>= 1;                   // three tiers x their classes must fit the class table
        if (const char *e = getenv("SVTK_POA_SUBCLASS")) { const int v = atoi(e); if (v == 1 || v == 2 || v == 4 || v == 8) sub = std::min(sub, v); }
        if (sub > 1)
            for (int s = 0; s < ns; ++s) {
                SetInfo &si = info[s];
                const double l2 = std::log2((double)std::max(si.maxlen, 1));
                si.cls = (int)std::ceil(l2 * sub - 1e-9);
            }
    }
    const long long cons_total = cons_off[ns] - cons_off[0];
    if (cons_off[0] != 0) return fail(SVTK_ERR_ARG, "cons_off must start at 0");

    // ---- device copies of the batch and of the outputs
    const size_t in_bytes = a16(8 * (ns + 1)) + a16(8 * (n_reads + 1)) + a16(n_bases) + a16(8 * (ns + 1));
    if (int rc = c->poa_in.reserve(in_bytes)) return rc;
    unsigned char *din = (unsigned char *)c->poa_in.p;
    long long *d_set_off = (long long *)din; din += a16(8 * (ns + 1));
    long long *d_seq_off = (long long *)din; din += a16(8 * (n_reads + 1));
    uint8_t *d_bases = din; din += a16(n_bases);
    long long *d_cons_off = (long long *)din;
    if (int rc = upload(c, d_set_off, b->set_off, 8 * (size_t)(ns + 1))) return rc;
    if (int rc = upload(c, d_seq_off, b->seq_off, 8 * (size_t)(n_reads + 1))) return rc;
    if (int rc = upload(c, d_bases, b->bases, (size_t)n_bases)) return rc;
    if (int rc = upload(c, d_cons_off, cons_off, 8 * (size_t)(ns + 1))) return rc;
    c->h2d_bytes += 16ull * (ns + 1) + 8ull * (n_reads + 1) + (uint64_t)n_bases;

    const size_t tr_bytes = trace ? (size_t)(a16(4 * n_reads) * 3 + a16(8 * n_reads) * 2) : 0;
    const size_t out_bytes = a16(cons_total) + (cons_cov ? a16(4 * cons_total) : 0) + a16(4 * ns) * 2 + tr_bytes + 64;
    if (int rc = c->poa_out.reserve(out_bytes)) return rc;
    unsigned char *dout = (unsigned char *)c->poa_out.p;
    uint8_t *d_cons = dout; dout += a16(cons_total);
    int32_t *d_cov = nullptr;
    if (cons_cov) { d_cov = (int32_t *)dout; dout += a16(4 * cons_total); }
    int32_t *d_len = (int32_t *)dout; dout += a16(4 * ns);
    int32_t *d_status = (int32_t *)dout; dout += a16(4 * ns);
    PoaTrace tr{};
    tr.dump_read = -1;
    if (trace) {
        tr.best_score = (int32_t *)dout; dout += a16(4 * n_reads);
        tr.node_n = (int32_t *)dout; dout += a16(4 * n_reads);
        tr.n_ops = (int32_t *)dout; dout += a16(4 * n_reads);
        tr.ops_hash = (uint64_t *)dout; dout += a16(8 * n_reads);
        tr.cells = (long long *)dout; dout += a16(8 * n_reads);
        CU(cudaMemsetAsync(tr.best_score, 0, tr_bytes, c->stream));
    }
    CU(cudaMemsetAsync(d_len, 0, a16(4 * ns) * 2, c->stream));

    // control block: cells[6] | 128 class cursors | 128 class descriptors | item list
    constexpr int MAXC = 256;
    const size_t ctl_bytes = 64 + 4 * MAXC + a16(sizeof(PoaClass) * MAXC) + 4 * (size_t)ns;
    if (int rc = c->poa_ctl.reserve(ctl_bytes)) return rc;
    unsigned long long *d_cells = (unsigned long long *)c->poa_ctl.p;
    int *d_cursors = (int *)((unsigned char *)c->poa_ctl.p + 64);
    PoaClass *d_classes = (PoaClass *)((unsigned char *)c->poa_ctl.p + 64 + 4 * MAXC);
    int *d_items = (int *)((unsigned char *)d_classes + a16(sizeof(PoaClass) * MAXC));
    for (int i = 0; i < 3; ++i) if (!c->poa_ev[i]) CU(cudaEventCreate(&c->poa_ev[i]));
    if (!c->poa_side) {
        CU(cudaStreamCreateWithFlags(&c->poa_side, cudaStreamNonBlocking));
        // First use: run both instantiations once with no work.  k_poa_sets<1> needs a larger local-memory pool than
        // k_poa_sets<8>; growing that pool is a device-wide synchronisation, which would otherwise serialise the first
        // pair of launches (measured: 12.1 s instead of 6.7 s for the first batch).
        CU(cudaMemsetAsync(d_classes, 0, sizeof(PoaClass) * MAXC, c->stream));
        PoaLaunch L0{};
        L0.classes = d_classes; L0.cursors = d_cursors; L0.scratch = (unsigned char *)c->poa_ctl.p;
        k_poa_sets<1><<<1, 32, 0, c->stream>>>(L0);
        k_poa_sets<8><<<1, 256, 0, c->stream>>>(L0);
        CU(cudaGetLastError());
        CU(cudaStreamSynchronize(c->stream));
    }

    size_t free_b = 0, total_b = 0;
    CU(cudaMemGetInfo(&free_b, &total_b));
    // Scratch is what limits how many heavy sets run side by side, and a batch with work for every worker ends when the work is
    // done: 10 000 C4-like sets run at 811 / 944 / 1 002 sets/s with 0.70 / 0.85 / 0.93 of the free memory.  Take 0.92 of it, and
    // make sure one block of that size can be had (fragmentation): probe, shrink by a tenth until it can.
    double budget_frac = 0.92;
    if (const char *e = getenv("SVTK_POA_BUDGET")) { const double v = atof(e); if (v > 0.05 && v <= 0.98) budget_frac = v; }
    long long budget = (long long)(budget_frac * (double)(free_b + c->poa_scratch.cap));
    for (int k = 0; k < 8 && budget > (long long)c->poa_scratch.cap; ++k) {
        void *probe = nullptr;
        if (cudaMalloc(&probe, (size_t)(budget - (long long)c->poa_scratch.cap)) == cudaSuccess) { cudaFree(probe); break; }
        (void)cudaGetLastError();
        budget = (long long)(0.9 * (double)budget);
    }

    std::vector<int> todo;
    for (int s = 0; s < ns; ++s) if (info[s].n > 0) todo.push_back(s);
    std::vector<int32_t> h_status(ns, 0);
    double scale = 1.0, kernel_ms = 0.0;
    if (const char *e = getenv("SVTK_POA_SCALE")) { const double v = atof(e); if (v > 0.0) scale = v; }   // tests: start from a too-small block
    int slots_per_sm = 32, slots_full = 8;            // warp-per-set CTAs asked for, per SM: with a wide tier / in a batch without one
    if (const char *e = getenv("SVTK_POA_SLOTS")) { const int v = atoi(e); if (v >= 1 && v <= 64) slots_per_sm = slots_full = v; }
    int wide_on = 1;
    if (const char *e = getenv("SVTK_POA_WIDE")) wide_on = atoi(e);                                        // experiments: 0 = one warp per set everywhere
    uint64_t relaunched = 0, max_workers = 0;
    CU(cudaMemsetAsync(d_cells, 0, 64, c->stream));
    int sms = 148;
    {
        int dev_sms = 0;
        if (cudaDeviceGetAttribute(&dev_sms, cudaDevAttrMultiProcessorCount, c->device) == cudaSuccess && dev_sms > 0) sms = dev_sms;
    }
    for (int attempt = 0; attempt < 6 && !todo.empty(); ++attempt) {
        // classes, largest reads first; inside a class heaviest set first (the atomic queue then balances the tail)
        std::sort(todo.begin(), todo.end(), [&](int x, int y) {
            if (info[x].cls != info[y].cls) return info[x].cls > info[y].cls;
            if (info[x].work != info[y].work) return info[x].work > info[y].work;
            return x < y;
        });
        // A set alone on a warp can outlast the rest of the batch: its reads are aligned one after the other, and only the row DP
        // (about half of its time) gets faster with more threads.  Two tiers by estimated work:
        //   k_poa_sets<8>  the sets within 1/1.5 of the heaviest one (reads beyond 2 kb; a 256-thread CTA holds more than half an
        //                  SM's registers) -- done in ~0.45 of their single-warp time;
        //   k_poa_sets<1>  everything else, one warp per set: the shape that fills the SMs.
        // (A third tier of 4-warp CTAs for the sets in between, and a common shared-memory carve-out so that all kernels share
        // SMs, were measured and dropped: DESIGN.md 6.)
        constexpr double WIDE_GAIN = 1.5;
        std::vector<int> wide, rest;
        {
            double wmax = 0;
            for (int x : todo) wmax = std::max(wmax, info[x].work);
            std::vector<int> by_work(todo);
            std::sort(by_work.begin(), by_work.end(), [&](int x, int y) { return info[x].work != info[y].work ? info[x].work > info[y].work : x < y; });
            std::vector<char> tier(ns, 0);
            // The wide tiers buy latency with registers and scratch; they pay only while the batch waits for its heaviest sets.
            // A batch with enough work to keep every warp-per-set worker busy for longer than its heaviest set takes on one warp
            // (r >= 1: total work / (workers x heaviest set)) gains nothing from them -- and a wide CTA keeps the warp-per-set
            // CTAs off its SM (different shared-memory carve-out), so with one on every SM nothing else starts for seconds
            // (10 000 sets: 22.7 s with 148 wide CTAs, the warp-per-set launch idle until 3.4 s; 11.1 s with none).  Measured
            // crossover on C4-like sets (batch time with / without the wide tier): 4 000 sets, r = 0.28: 6.7 / 7.5 s; 5 000,
            // r = 0.35: 7.5 / 7.9 s; 6 000, r = 0.42: 8.5 / 7.8 s.  No wide tier from r = 0.39 on, and never more wide CTAs
            // than 5/6 of the SMs.
            int wide_cap = std::max(1, sms * 5 / 6);
            {
                std::vector<PoaClass> cls0;
                long long w0 = 0, m0 = 0;
                double tot = 0;
                for (int x : todo) tot += info[x].work;
                if (plan_group(info, todo.data(), todo.size(), 0, scale, (long long)sms * slots_per_sm, budget, cls0, w0, m0) == 0 && w0 > 0) {
                    // workers that count: the machine's set rate stops growing at about 8 warp-per-set workers per SM (10 000 sets:
                    // 990 sets/s with 1 150 workers, 970 with 2 246 -- a third of the warp cycles then wait for instruction fetch)
                    const double r = tot / ((double)std::min<long long>(w0, (long long)sms * 8) * std::max(wmax, 1.0));
                    if (r >= 0.39) wide_cap = 0;
                    if (getenv("SVTK_POA_DEBUG")) fprintf(stderr, "[poa] warp-per-set plan alone: %lld workers, r = %.2f -> at most %d wide CTAs\n", w0, r, wide_cap);
                }
            }
            for (int x : by_work) {
                if (wide_on < 1 || info[x].maxlen <= 2048) break;
                if ((int)wide.size() < wide_cap && info[x].work * WIDE_GAIN >= wmax) { wide.push_back(x); tier[x] = 2; }
                else break;
            }
            for (int x : todo) if (!tier[x]) rest.push_back(x);
            auto by_class = [&](int x, int y) {
                if (info[x].cls != info[y].cls) return info[x].cls > info[y].cls;
                if (info[x].work != info[y].work) return info[x].work > info[y].work;
                return x < y;
            };
            std::sort(wide.begin(), wide.end(), by_class);
            todo = wide;
            todo.insert(todo.end(), rest.begin(), rest.end());
        }
        const size_t nA = wide.size();
        std::vector<PoaClass> clsA, clsB;
        long long wA = 0, wB = 0, memA = 0, memB = 0;
        if (int rc = plan_group(info, todo.data(), nA, 0, scale, (long long)nA, (long long)(0.45 * (double)budget), clsA, wA, memA)) return rc;
        // A batch without a wide tier has work for every worker: what counts is the set rate of the machine, and that peaks at 8
        // warp-per-set CTAs per SM (10 000 C4-like sets, sets/s at 6 / 7 / 8 / 10 / 12 / 15 per SM: 854 / 970 / 1 061 / 989 / 944 /
        // 955 -- the kernel is ~150 KB of SASS and every warp is in another phase: beyond two warps per scheduler a third of the
        // warp cycles wait for instruction fetch, profiles/r03y).  With a wide tier the batch is waiting for its heaviest sets and
        // the small ones are better out of the way early (4 000 sets: 6.4 s uncapped, 6.8 s capped).
        const long long narrow_W = (long long)sms * (nA == 0 ? slots_full : slots_per_sm);
        if (int rc = plan_group(info, todo.data() + nA, todo.size() - nA, (int)nA, scale, narrow_W, budget - memA, clsB, wB, memB)) return rc;
        if (clsA.size() + clsB.size() > (size_t)MAXC) return fail(SVTK_ERR_ARG, "poa: more than %d size classes", MAXC);
        for (PoaClass &pc : clsB) pc.scratch_off += memA;
        std::vector<PoaClass> cls(clsA);
        cls.insert(cls.end(), clsB.begin(), clsB.end());
        if (int rc = c->poa_scratch.reserve((size_t)(memA + memB))) return rc;
        CU(cudaMemcpyAsync(d_items, todo.data(), 4 * todo.size(), cudaMemcpyHostToDevice, c->stream));
        CU(cudaMemcpyAsync(d_classes, cls.data(), sizeof(PoaClass) * cls.size(), cudaMemcpyHostToDevice, c->stream));
        CU(cudaMemsetAsync(d_cursors, 0, 4 * MAXC, c->stream));
        PoaLaunch L{};
        L.items = d_items;
        L.set_off = d_set_off; L.seq_off = d_seq_off; L.bases = d_bases; L.cons = d_cons; L.cov = d_cov; L.cons_off = d_cons_off;
        L.cons_len = d_len; L.status = d_status; L.cells = d_cells; L.scratch = (unsigned char *)c->poa_scratch.p;
        L.trace = tr; L.have_trace = trace ? 1 : 0;
        const bool dbg = getenv("SVTK_POA_DEBUG") != nullptr;
        if (dbg) {
            if (int rc = c->tmp_c.reserve(32 * (size_t)ns + 16)) return rc;
            CU(cudaMemsetAsync(c->tmp_c.p, 0, 32 * (size_t)ns, c->stream));
            L.times = (unsigned long long *)c->tmp_c.p; L.n_sets_dbg = ns;
        }
        CU(cudaEventRecord(c->poa_ev[0], c->stream));
        if (wA > 0) {                                          // side streams: the three launches are resident together
            CU(cudaStreamWaitEvent(c->poa_side, c->poa_ev[0], 0));
            PoaLaunch LA = L;
            LA.n_classes = (int)clsA.size(); LA.classes = d_classes; LA.cursors = d_cursors;
            k_poa_sets<8><<<(unsigned)wA, 256, 0, c->poa_side>>>(LA);
            CU(cudaGetLastError());
            CU(cudaEventRecord(c->poa_ev[2], c->poa_side));
            c->launches += 1;
        }
        if (wB > 0) {
            PoaLaunch LB = L;
            LB.n_classes = (int)clsB.size(); LB.classes = d_classes + clsA.size(); LB.cursors = d_cursors + clsA.size();
            k_poa_sets<1><<<(unsigned)wB, 32, 0, c->stream>>>(LB);
            CU(cudaGetLastError());
            c->launches += 1;
        }
        if (wA > 0) CU(cudaStreamWaitEvent(c->stream, c->poa_ev[2], 0));
        CU(cudaEventRecord(c->poa_ev[1], c->stream));
        CU(cudaMemcpyAsync(h_status.data(), d_status, 4 * (size_t)ns, cudaMemcpyDeviceToHost, c->stream));
        CU(cudaStreamSynchronize(c->stream));                  // also: todo / cls were read by the copies above
        float ms = 0.f;
        CU(cudaEventElapsedTime(&ms, c->poa_ev[0], c->poa_ev[1]));
        kernel_ms += ms;
        if (getenv("SVTK_POA_DEBUG")) {
            float msA = 0.f;
            if (wA > 0) cudaEventElapsedTime(&msA, c->poa_ev[0], c->poa_ev[2]);
            std::vector<unsigned long long> tm(4 * (size_t)ns);
            cudaMemcpy(tm.data(), c->tmp_c.p, 32 * (size_t)ns, cudaMemcpyDeviceToHost);
            unsigned long long t0 = ~0ull;
            for (int x : todo) if (tm[2 * x]) t0 = std::min(t0, tm[2 * x]);
            for (size_t k = 0; k < cls.size(); ++k) {
                double first = 1e30, last = 0, sum = 0, lstart = 0, use_cells = 0, use_nodes = 0, use_edges = 0;
                for (int i = cls[k].item_begin; i < cls[k].item_end; ++i) {
                    const int x = todo[i];
                    {   // how much of the class's block its sets needed (cells: 5 planes x 2 or 4 bytes -- 4 assumed, the upper bound)
                        const PoaCaps own = caps_for(info[x], scale);
                        use_cells = std::max(use_cells, (double)tm[2 * ns + 2 * x] * (4.0 * NPL) / (double)std::max<long long>(own.cell_bytes, 1));
                        use_nodes = std::max(use_nodes, (double)(tm[2 * ns + 2 * x + 1] & 0xffffffffull) / (double)own.node_cap);
                        use_edges = std::max(use_edges, (double)(tm[2 * ns + 2 * x + 1] >> 32) / (double)own.edge_cap);
                    }
                    const double a = (tm[2 * x] - t0) * 1e-6, b = (tm[2 * x + 1] - t0) * 1e-6;
                    first = std::min(first, a); last = std::max(last, b); lstart = std::max(lstart, a); sum += b - a;
                }
                fprintf(stderr, "[poa]   class %zu %s cls %d: %d sets, %d workers, %.1f MB/worker; first start %.0f ms, last start %.0f ms, last end %.0f ms, busy %.1f worker-s; most of a set's own estimate used: cells %.2f nodes %.2f edges %.2f\n",
                        k, k < clsA.size() ? "wide" : "warp", info[todo[cls[k].item_begin]].cls, cls[k].item_end - cls[k].item_begin,
                        cls[k].worker_end - cls[k].worker_begin, poa_layout(cls[k].caps).total / 1e6, first, lstart, last, sum * 1e-3, use_cells, use_nodes, use_edges);
            }
            std::vector<int> by_end(todo);
            std::sort(by_end.begin(), by_end.end(), [&](int x, int y) { return tm[2 * x + 1] > tm[2 * y + 1]; });
            for (size_t i = 0; i < std::min<size_t>(12, by_end.size()); ++i) {
                const int x = by_end[i];
                fprintf(stderr, "[poa]   late set %d: cls %d maxlen %d reads %d work %.3g: %.0f -> %.0f ms\n", x, info[x].cls, info[x].maxlen, info[x].n, info[x].work,
                        (tm[2 * x] - t0) * 1e-6, (tm[2 * x + 1] - t0) * 1e-6);
            }
            fprintf(stderr, "[poa] attempt %d: <8> %lld sets / %lld CTAs (%.1f MB) done at %.1f ms; warp-per-set %zu sets / %lld CTAs (%.1f MB); all done at %.1f ms\n",
                    attempt, (long long)nA, wA, memA / 1e6, msA, todo.size() - nA, wB, memB / 1e6, ms);
        }
        max_workers = std::max<uint64_t>(max_workers, (uint64_t)(wA + wB));
        std::vector<int> again;
        for (int s : todo) {
            const int st = h_status[s];
            if (st == POA_ERR_NODES || st == POA_ERR_EDGES || st == POA_ERR_CELLS || st == POA_ERR_CIGAR) again.push_back(s);
        }
        relaunched += again.size();
        todo.swap(again);
        scale *= 4.0;
    }

    // ---- results back
    std::vector<int32_t> h_len(ns);
    CU(cudaMemcpyAsync(h_len.data(), d_len, 4 * (size_t)ns, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaMemcpyAsync(cons, d_cons, (size_t)cons_total, cudaMemcpyDeviceToHost, c->stream));
    if (cons_cov) CU(cudaMemcpyAsync(cons_cov, d_cov, 4 * (size_t)cons_total, cudaMemcpyDeviceToHost, c->stream));
    if (trace) {
        CU(cudaMemcpyAsync(trace->best_score, tr.best_score, 4 * (size_t)n_reads, cudaMemcpyDeviceToHost, c->stream));
        CU(cudaMemcpyAsync(trace->node_n, tr.node_n, 4 * (size_t)n_reads, cudaMemcpyDeviceToHost, c->stream));
        CU(cudaMemcpyAsync(trace->n_ops, tr.n_ops, 4 * (size_t)n_reads, cudaMemcpyDeviceToHost, c->stream));
        CU(cudaMemcpyAsync(trace->ops_hash, tr.ops_hash, 8 * (size_t)n_reads, cudaMemcpyDeviceToHost, c->stream));
        CU(cudaMemcpyAsync(trace->cells, tr.cells, 8 * (size_t)n_reads, cudaMemcpyDeviceToHost, c->stream));
    }
    unsigned long long h_cells[6] = {0, 0, 0, 0, 0, 0};
    CU(cudaMemcpyAsync(h_cells, d_cells, 48, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    c->d2h_bytes += (uint64_t)cons_total * (cons_cov ? 5 : 1) + 8ull * ns;
    int n_failed = 0, first_failed = -1, first_code = 0;
    for (int s = 0; s < ns; ++s) {
        const bool bad = h_status[s] != POA_OK;
        cons_len[s] = bad ? 0 : h_len[s];
        if (set_status) set_status[s] = bad ? SVTK_POA_FAILED : SVTK_POA_OK;
        if (bad) { if (!n_failed) { first_failed = s; first_code = h_status[s]; } ++n_failed; }
    }
    if (stats) {
        stats->n_sets = (uint64_t)ns; stats->n_reads = (uint64_t)n_reads; stats->n_bases = (uint64_t)n_bases;
        stats->cells = h_cells[0]; stats->cells_full = h_cells[1]; stats->relaunched = relaunched; stats->workers = max_workers;
        stats->kernel_ms = kernel_ms;
        for (int k = 0; k < 4; ++k) stats->phase_cycles[k] = h_cells[2 + k];
    }
    if (n_failed) return fail(SVTK_ERR_STATE, "poa: %d read set(s) failed, first set %d with internal code %d", n_failed, first_failed, first_code);
    return 0;
}
