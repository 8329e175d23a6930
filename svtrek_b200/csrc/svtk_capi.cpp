// svtk_capi.cpp -- the C ABI of include/svtk.h: context, HBM residency of the read set, pinned staging,
// stream-ordered launches, overflow resolution.  Host C++ only; kernels live in svtk_kernels.cu.
#include "../../include/svtk.h"
#include "svtk_ctx.h"

#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

namespace svtk {

static thread_local std::string g_err;

int fail(int code, const char *fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_err = buf;
    return code;
}

#define CU(call)                                                                                              \
    do {                                                                                                      \
        cudaError_t e_ = (call);                                                                              \
        if (e_ != cudaSuccess) return fail(SVTK_ERR_CUDA, "%s: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

// ---- small RAII-less buffer helpers (the ctx owns everything and frees in svtk_destroy)
int DevBuf::reserve(size_t bytes) {
    if (bytes <= cap) return 0;
    if (p) cudaFree(p);
    p = nullptr; cap = 0;
    size_t want = std::max(bytes, (size_t)256);
    want = (want + 255) & ~(size_t)255;
    cudaError_t e = cudaMalloc(&p, want);
    if (e != cudaSuccess) return fail(SVTK_ERR_ALLOC, "cudaMalloc(%zu): %s", want, cudaGetErrorString(e));
    cap = want;
    return 0;
}
void DevBuf::release() { if (p) cudaFree(p); p = nullptr; cap = 0; }

int PinBuf::reserve(size_t bytes) {
    if (bytes <= cap) return 0;
    if (p) cudaFreeHost(p);
    p = nullptr; cap = 0;
    size_t want = std::max(bytes, (size_t)4096);
    cudaError_t e = cudaMallocHost(&p, want);
    if (e != cudaSuccess) return fail(SVTK_ERR_ALLOC, "cudaMallocHost(%zu): %s", want, cudaGetErrorString(e));
    cap = want;
    return 0;
}
void PinBuf::release() { if (p) cudaFreeHost(p); p = nullptr; cap = 0; }

// host (pageable) -> device through two pinned staging buffers so the CPU copy of chunk k+1 overlaps the DMA of chunk k
int upload(Ctx *c, void *dst, const void *src, size_t bytes) {
    if (bytes == 0) return 0;
    {   // caller's buffer already page-locked (cudaHostAlloc / cudaHostRegister): DMA straight from it
        cudaPointerAttributes at;
        if (cudaPointerGetAttributes(&at, src) == cudaSuccess && at.type == cudaMemoryTypeHost) {
            CU(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, c->stream));
            return 0;
        }
        cudaGetLastError();
    }
    const size_t CH = 32u << 20;
    for (int i = 0; i < 2; ++i) {
        if (int rc = c->stage[i].reserve(std::min(bytes, CH))) return rc;
        if (!c->stage_ev[i]) CU(cudaEventCreateWithFlags(&c->stage_ev[i], cudaEventDisableTiming));
    }
    size_t done = 0;
    int k = 0;
    while (done < bytes) {
        const size_t n = std::min(CH, bytes - done);
        CU(cudaEventSynchronize(c->stage_ev[k]));
        memcpy(c->stage[k].p, (const char *)src + done, n);
        CU(cudaMemcpyAsync((char *)dst + done, c->stage[k].p, n, cudaMemcpyHostToDevice, c->stream));
        CU(cudaEventRecord(c->stage_ev[k], c->stream));
        done += n;
        k ^= 1;
    }
    return 0;
}

int set_device(Ctx *c) {
    CU(cudaSetDevice(c->device));
    return 0;
}

static int finish_reads(Ctx *c) {
    c->reads_from_bam = false;            // svtk_load_bam sets it again after adopting its own arrays
    c->reads.tiles = nullptr; c->reads.tile_off = nullptr;
    // the svtk_reads contract, checked on the host where it is cheap: tid_off monotone inside [0, n_reads]
    for (size_t t = 0; t + 1 < c->h_tid_off.size(); ++t)
        if (c->h_tid_off[t] < 0 || c->h_tid_off[t] > c->h_tid_off[t + 1] || c->h_tid_off[t + 1] > c->reads.n_reads)
            return fail(SVTK_ERR_ARG, "svtk_reads: tid_off must be non-decreasing within [0, n_reads] (reference %zu)", t);
    if (!c->h_tid_off.empty() && c->h_tid_off[0] != 0) return fail(SVTK_ERR_ARG, "svtk_reads: tid_off[0] must be 0");
    // planner look-back bound: max(end - pos) over the read set; the packed-CIGAR / coordinate-order contract; the last read
    // position of every reference (sizes the position tiles)
    const int n_tid = c->reads.n_tid;
    if (int rc = c->scratch4.reserve(64 + 4 * (size_t)std::max(n_tid, 1))) return rc;
    CU(cudaMemsetAsync(c->scratch4.p, 0, 8, c->stream));
    CU(launch_max_span(c->reads.pos, c->reads.end, c->reads.n_reads, (uint32_t *)c->scratch4.p, c->stream));
    CU(launch_check_layout(c->reads, (unsigned long long)c->n_words, (unsigned int *)c->scratch4.p + 1, c->stream));
    CU(launch_last_pos(c->reads, (uint32_t *)c->scratch4.p + 16, c->stream));
    c->launches += 3 * (c->reads.n_reads > 0);
    std::vector<uint32_t> back(16 + (size_t)std::max(n_tid, 1), 0u);
    back[0] = 1;
    CU(cudaMemcpyAsync(back.data(), c->scratch4.p, 64 + 4 * (size_t)n_tid, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    if (c->reads.n_reads > 0) {
        if (back[1] & 1u) return fail(SVTK_ERR_ARG, "svtk_reads: CIGARs must be packed back to back in read order (cigar_off[r] + n_cigar[r] == cigar_off[r+1], cigar_off[0] == 0, last end == n_words)");
        if (back[1] & 2u) return fail(SVTK_ERR_ARG, "svtk_reads: reads must be sorted by pos within each reference (BAM coordinate order)");
        if (back[1] & 4u) return fail(SVTK_ERR_ARG, "svtk_reads: end must be >= pos (bam_endpos)");
    }
    c->reads.max_span = std::max(back[0], 1u);
    // position tiles: per reference, the first read at or right of every 4 096-base boundary -- what the .bai's linear index is to
    // the reference's iterator (hts.c:3099-3111).  Built once per load; k_refine_warp starts its two read-range searches there.
    if (c->reads.n_reads > 0 && n_tid > 0) {
        std::vector<long long> toff((size_t)n_tid + 1, 0);
        for (int t = 0; t < n_tid; ++t) {
            const bool has = c->h_tid_off[(size_t)t + 1] > c->h_tid_off[(size_t)t];
            toff[(size_t)t + 1] = toff[(size_t)t] + (has ? (long long)(back[16 + (size_t)t] >> pos_tile_shift()) + 2 : 0);
        }
        if (int rc = c->pos_tiles.reserve(4 * (size_t)std::max(toff.back(), 1ll))) return rc;
        if (int rc = c->pos_tile_off.reserve(8 * ((size_t)n_tid + 1))) return rc;
        if (int rc = upload(c, c->pos_tile_off.p, toff.data(), 8 * ((size_t)n_tid + 1))) return rc;
        CU(launch_pos_tiles(c->reads, (uint32_t *)c->pos_tiles.p, (const long long *)c->pos_tile_off.p, c->stream));
        c->launches += 1;
        CU(cudaStreamSynchronize(c->stream));               // `toff` is pageable: the upload above staged it, but keep the contract simple
        c->reads.tiles = (const uint32_t *)c->pos_tiles.p; c->reads.tile_off = (const long long *)c->pos_tile_off.p;
    }
    c->have_reads = true;
    c->index_valid = false;
    return 0;
}

// ---- control block: the index counters followed by the batch flags, so that a step which builds the index clears both with
// one memset and reads both back with one copy.  flags: [0] overflow count  [1] fail flag  [2..3] arena top (u64)
static size_t ctl_index_bytes() { return 4 * (size_t)index_ctrl_words(); }
// Two blocks on the device: an index build flips to the other one, and the index kernel itself clears the block the NEXT build
// will use (index_zeroes_next_ctrl), so a steady stream of builds issues no memset.  ctl_clean[b]: block b is known to be zero.
static size_t ctl_block_bytes() { return (ctl_index_bytes() + 64 + 255) & ~(size_t)255; }
static int ensure_ctl(Ctx *c) {
    const void *before = c->idx_ctrl.p;
    if (int rc = c->idx_ctrl.reserve(2 * ctl_block_bytes())) return rc;
    if (c->idx_ctrl.p != before) c->ctl_clean[0] = c->ctl_clean[1] = false;
    return c->h_idx.reserve(ctl_index_bytes() + 64);
}
static char *ctl_base(Ctx *c, int which) { return (char *)c->idx_ctrl.p + (size_t)which * ctl_block_bytes(); }
static int *ov_flags(Ctx *c) { return (int *)(ctl_base(c, c->ctl_cur) + ctl_index_bytes()); }
static int *host_flags(Ctx *c) { return (int *)((char *)c->h_idx.p + ctl_index_bytes()); }

// ---- signature index (svtk_index.cu)
static int ensure_index_buffers(Ctx *c, uint64_t pool_cap, uint64_t seg_cap) {
    const long long nr = c->reads.n_reads;
    const int per = index_reads_per_chunk();
    const long long n_chunks = (nr + per - 1) / per;
    const uint64_t ev_cap = pool_cap * (uint64_t)index_pools() + (uint64_t)n_chunks * (uint64_t)index_fixed_slots();
    if (ev_cap >= (1ull << 31) || seg_cap >= (1ull << 31)) return fail(SVTK_ERR_ARG, "signature index beyond 2^31 events: shard the read set");
    if (int rc = c->idx_events.reserve(16 * (size_t)ev_cap)) return rc;
    if (int rc = c->idx_chunks.reserve(16 * (size_t)std::max(n_chunks, 1ll))) return rc;
    if (int rc = c->idx_segs.reserve(16 * (size_t)seg_cap)) return rc;
    if (int rc = c->idx_rend.reserve(4 * (size_t)std::max(nr, 1ll))) return rc;
    if (int rc = ensure_ctl(c)) return rc;
    c->idx = LaunchIndex{c->idx_events.p, c->idx_chunks.p, c->idx_segs.p, (uint32_t *)c->idx_rend.p, (unsigned int *)ctl_base(c, c->ctl_cur),
                         (unsigned int)pool_cap, (unsigned int)seg_cap, (int)n_chunks, nullptr, 0};
    return 0;
}

// harvest finished kernel timings.  which: 1 = refine pair, 2 = index pair, 3 = both.  A pair is only waited for when it is
// about to be re-recorded (or read); svtk_sync harvests both after the stream has drained, so a step that ends in svtk_sync
// never blocks the host between its two kernels.
static int collect_timing(Ctx *c, int which = 3) {
    if ((which & 1) && c->prof_pending) {
        c->prof_pending = false;
        CU(cudaEventSynchronize(c->prof_ev[1]));
        float ms = 0.f;
        CU(cudaEventElapsedTime(&ms, c->prof_ev[0], c->prof_ev[1]));
        c->refine_ms += (double)ms;
        c->refine_launches += 1;
    }
    if ((which & 2) && c->prof_idx_pending) {
        c->prof_idx_pending = false;
        CU(cudaEventSynchronize(c->prof_ev[3]));
        float ms = 0.f;
        CU(cudaEventElapsedTime(&ms, c->prof_ev[2], c->prof_ev[3]));
        c->index_ms += (double)ms;
        c->index_launches += 1;
    }
    return 0;
}

static int build_index(Ctx *c) {
    const int per = index_reads_per_chunk();
    if (c->idx.pool_cap == 0 || c->idx.n_chunks != (int)((c->reads.n_reads + per - 1) / per)) {
        // first guess: one event per 16 CIGAR words plus both clip events of every read, spread over the sub-pools with
        // 25% slack for imbalance; grown on demand (resolve_overflow)
        const uint64_t ev = (uint64_t)c->n_words / 16 + 2 * (uint64_t)c->reads.n_reads;
        const uint64_t pool = ev / (uint64_t)index_pools() * 5 / 4 + 1024;
        if (int rc = ensure_index_buffers(c, std::max<uint64_t>(pool, c->idx.pool_cap), std::max<uint64_t>(ev / 32 + 1024, c->idx.seg_cap))) return rc;
    }
    // counters and batch flags of this build live in the other control block: already cleared by the previous build's kernel
    // (steady state), else by a memset here
    const bool kz = index_zeroes_next_ctrl() != 0 && c->reads.n_reads > 0;
    c->ctl_cur ^= 1;
    if (!c->ctl_clean[c->ctl_cur]) CU(cudaMemsetAsync(ctl_base(c, c->ctl_cur), 0, ctl_index_bytes() + 64, c->stream));
    c->idx.ctrl = (unsigned int *)ctl_base(c, c->ctl_cur);
    c->idx.next_ctrl = kz ? (unsigned int *)ctl_base(c, c->ctl_cur ^ 1) : nullptr;
    c->idx.next_words = (int)((ctl_index_bytes() + 64) / 4);
    c->ctl_clean[c->ctl_cur] = false;
    c->ctl_clean[c->ctl_cur ^ 1] = kz;
    c->flags_clean = true;
    const bool prof = c->profile && (c->prof_tick_idx++ % (unsigned long long)c->prof_stride) == 0;
    if (prof) { if (int rc = collect_timing(c, 2)) return rc; CU(cudaEventRecord(c->prof_ev[2], c->stream)); }
    CU(launch_index_reads(c->reads, c->idx, c->stream));
    if (prof) { CU(cudaEventRecord(c->prof_ev[3], c->stream)); c->prof_idx_pending = true; }
    c->launches += c->reads.n_reads > 0;
    c->index_valid = true;
    c->index_fresh = true;
    return 0;
}

static int ensure_batch(Ctx *c, int n_win) {
    if (int rc = c->range.reserve((size_t)n_win * sizeof(int2))) return rc;
    if (int rc = c->ov_items.reserve((size_t)n_win * sizeof(int2))) return rc;
    return ensure_ctl(c);
}

enum { MODE_INDEXED = 0, MODE_WALKER_STATS = 1, MODE_INDEXED_STATS = 2 };

static int launch_batch(Ctx *c, const LaunchWindows &w, int32_t *dev_result, int mode, char *host_ctl = nullptr) {
    if (!c->have_reads) return fail(SVTK_ERR_STATE, "no reads loaded (svtk_load_reads / svtk_attach_reads_device)");
    if (int rc = set_device(c)) return rc;
    if (int rc = ensure_batch(c, std::max(w.n_win, 1))) return rc;
    c->pending = false;
    if (w.n_win <= 0) return 0;
    const bool with_index = mode != MODE_WALKER_STATS && !c->index_valid;      // build_index clears the whole control block
    // an index built by an explicit svtk_index_reads since the last batch cleared the flags with its counters
    if (!with_index && !c->flags_clean) CU(cudaMemsetAsync(ov_flags(c), 0, 64, c->stream));
    unsigned long long *counters = nullptr;
    if (mode != MODE_INDEXED) {
        if (int rc = c->counters.reserve(64)) return rc;
        CU(cudaMemsetAsync(c->counters.p, 0, 64, c->stream));
        counters = (unsigned long long *)c->counters.p;
    }
    if (mode == MODE_WALKER_STATS) {
        // the reference-shaped walker (one CIGAR walk per (window, read) pair): instrumented cross-check
        CU(launch_plan(c->reads, w, (int2 *)c->range.p, c->stream));
        CU(launch_refine_fused(c->reads, w, (const int2 *)c->range.p, c->params.consensus_interval_range, c->params.consensus_interval,
                               c->params.consensus_min_count, dev_result, ov_flags(c), (int2 *)c->ov_items.p, w.n_win, counters,
                               c->stream));
        c->launches += 2;
    } else {
        if (!c->index_valid) if (int rc = build_index(c)) return rc;
        const bool prof = c->profile && (c->prof_tick_ref++ % (unsigned long long)c->prof_stride) == 0;
        if (prof) {
            if (int rc = collect_timing(c, 1)) return rc;
            CU(cudaEventRecord(c->prof_ev[0], c->stream));
        }
        CU(launch_refine_indexed(c->reads, w, c->idx, (int2 *)c->range.p, c->params.consensus_interval_range, c->params.consensus_interval,
                                 c->params.consensus_min_count, dev_result, ov_flags(c), (int2 *)c->ov_items.p, w.n_win, counters,
                                 c->stream));
        if (prof) {
            CU(cudaEventRecord(c->prof_ev[1], c->stream));
            c->prof_pending = true;
        }
        c->launches += 1;
    }
    if (host_ctl)            // svtk_audit_device: the whole control block of this step goes to the step's own slot
        CU(cudaMemcpyAsync(host_ctl, ctl_base(c, c->ctl_cur), ctl_index_bytes() + 4, cudaMemcpyDeviceToHost, c->stream));
    else if (mode != MODE_WALKER_STATS && c->index_fresh)
        CU(cudaMemcpyAsync(c->h_idx.p, ctl_base(c, c->ctl_cur), ctl_index_bytes() + 4, cudaMemcpyDeviceToHost, c->stream));
    else
        CU(cudaMemcpyAsync(host_flags(c), ov_flags(c), 4, cudaMemcpyDeviceToHost, c->stream));
    c->flags_clean = false;                                   // this batch's kernels own the flags now (a build inside it set it)
    c->pending = true;
    c->pend_mode = mode;
    c->pend_win = w;
    c->pend_result = dev_result;
    return 0;
}

// after the stream drained: (1) an index built by this batch that ran out of room is rebuilt at its exact size and the
// batch is redone; (2) windows that overflowed the fast path's shared segment are redone by the overflow kernel
// totals of the index control block the last fresh build copied to the host
static void index_totals(const Ctx *c, uint64_t *events, uint64_t *segs, unsigned int *max_pool, unsigned int *err) {
    const unsigned int *ctrl = (const unsigned int *)c->h_idx.p;
    const int st = index_ctrl_stride(), np = index_pools();
    uint64_t ev = 0; unsigned int mx = 0;
    for (int p = 0; p < np; ++p) { ev += ctrl[p * st]; mx = std::max(mx, ctrl[p * st]); }
    if (index_fixed_slots()) for (int p = 0; p < np; ++p) ev += ctrl[p * st + 1];          // events in the chunks' own slots
    *events = ev; *segs = ctrl[np * st]; *max_pool = mx; *err = ctrl[(np + 1) * st];
}

static int resolve_overflow(Ctx *c) {
    if (!c->pending) return 0;
    c->pending = false;
    if (c->index_fresh) {
        c->index_fresh = false;
        unsigned int max_pool = 0, err = 0;
        index_totals(c, &c->idx_n_events, &c->idx_n_segs, &max_pool, &err);
        if (err) {
            const LaunchWindows w = c->pend_win;
            int32_t *res = c->pend_result;
            const int mode = c->pend_mode;
            // every flush bumped its counter even when it found no room, so the tops are what THIS layout needed; a rebuild can
            // need a little more (a chunk whose earlier flushes failed kept its last events in its own slots and never counted
            // them): grow with slack and try again until the build goes through
            uint64_t pool_need = (uint64_t)max_pool + 64, seg_need = c->idx_n_segs + c->idx_n_events / 32 + 1024;
            for (int attempt = 0;; ++attempt) {
                if (int rc = ensure_index_buffers(c, pool_need, seg_need)) return rc;
                c->index_valid = false;
                if (int rc = launch_batch(c, w, res, mode)) return rc;
                CU(cudaStreamSynchronize(c->stream));
                unsigned int err2 = 0; uint64_t e2 = 0, s2 = 0;
                index_totals(c, &e2, &s2, &max_pool, &err2);
                if (!err2) break;
                if (attempt >= 6) return fail(SVTK_ERR_STATE, "signature index still out of room after re-dimensioning (events %llu, segments %llu)",
                                              (unsigned long long)e2, (unsigned long long)s2);
                pool_need = std::max<uint64_t>(pool_need * 3 / 2, (uint64_t)max_pool + 4096);
                seg_need = std::max<uint64_t>(seg_need * 3 / 2, s2 + 4096);
            }
            return resolve_overflow(c);
        }
    }
    const int n_ov = *host_flags(c);
    c->last_overflow = n_ov;
    if (n_ov <= 0) return 0;
    std::vector<int2> items((size_t)n_ov);
    CU(cudaMemcpyAsync(items.data(), c->ov_items.p, sizeof(int2) * (size_t)n_ov, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    const bool walker = c->pend_mode == MODE_WALKER_STATS;
    unsigned long long arena_need = 0;
    const int big = walker ? big_segment_ints() : refine_block_segment_ints();
    for (const int2 &it : items) {
        unsigned long long p = 1; while (p < (unsigned long long)it.y) p <<= 1;
        if (p > (unsigned long long)big) arena_need += p;
    }
    if (arena_need) if (int rc = c->arena.reserve(arena_need * 4)) return rc;
    int *flags = ov_flags(c);
    CU(cudaMemsetAsync(flags + 1, 0, 12, c->stream));     // fail flag + arena top
    if (walker)
        CU(launch_refine_overflow(c->reads, c->pend_win, (const int2 *)c->range.p, c->params.consensus_interval_range,
                                  c->params.consensus_interval, c->params.consensus_min_count, c->pend_result, flags, (int2 *)c->ov_items.p,
                                  c->pend_win.n_win, n_ov, (int *)c->arena.p, arena_need, (unsigned long long *)(flags + 2), flags + 1,
                                  c->stream));
    else
        CU(launch_refine_block(c->reads, c->pend_win, c->idx, (const int2 *)c->range.p, c->params.consensus_interval_range,
                               c->params.consensus_interval, c->params.consensus_min_count, c->pend_result, flags, (int2 *)c->ov_items.p,
                               c->pend_win.n_win, n_ov, (int *)c->arena.p, arena_need, (unsigned long long *)(flags + 2), flags + 1,
                               c->stream));
    c->launches += 1;
    CU(cudaMemcpyAsync(host_flags(c) + 1, flags + 1, 4, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    if (host_flags(c)[1]) return fail(SVTK_ERR_STATE, "overflow arena exhausted");
    return 0;
}

static int check_reads(const svtk_reads *r) {
    if (!r || r->n_tid < 0 || r->n_reads < 0 || r->n_words < 0) return fail(SVTK_ERR_ARG, "bad svtk_reads");
    if (r->n_reads >= (1ll << 31) - 1024) return fail(SVTK_ERR_ARG, "more than 2^31 reads per device: shard the read set");
    if (r->n_tid > 0 && !r->tid_off) return fail(SVTK_ERR_ARG, "tid_off is NULL");
    if (r->n_reads > 0 && (!r->pos || !r->end || !r->n_cigar || !r->cigar_off)) return fail(SVTK_ERR_ARG, "read arrays are NULL");
    if (r->n_words > 0 && !r->cigar) return fail(SVTK_ERR_ARG, "cigar is NULL");
    return 0;
}

int adopt_device_reads(Ctx *c, const svtk_reads *d) {
    if (int rc = check_reads(d)) return rc;
    if (((uintptr_t)d->cigar & 15) != 0) return fail(SVTK_ERR_ARG, "device cigar array must be 16-byte aligned");
    if (int rc = set_device(c)) return rc;
    c->have_reads = false;
    if (int rc = c->r_tid_off.reserve(sizeof(long long) * (size_t)(d->n_tid + 1))) return rc;
    c->h_tid_off.assign(d->n_tid + 1, 0);
    if (d->n_tid > 0) memcpy(c->h_tid_off.data(), d->tid_off, sizeof(int64_t) * (size_t)(d->n_tid + 1));
    if (c->h_tid_off.back() != d->n_reads) return fail(SVTK_ERR_ARG, "tid_off[n_tid] != n_reads");
    if (int rc = upload(c, c->r_tid_off.p, c->h_tid_off.data(), sizeof(long long) * (size_t)(d->n_tid + 1))) return rc;
    c->reads = LaunchReads{d->n_tid, d->n_reads, (const long long *)c->r_tid_off.p, d->pos, d->end, d->n_cigar,
                           (const unsigned long long *)d->cigar_off, d->cigar, 1, nullptr, nullptr};
    c->n_words = d->n_words;
    return finish_reads(c);
}

}  // namespace svtk

using namespace svtk;

extern "C" {

int svtk_version(void) { return SVTK_VERSION; }
const char *svtk_last_error(void) { return g_err.c_str(); }

int svtk_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

svtk_ctx *svtk_create(int device, const svtk_params *params, void *cuda_stream) {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n <= 0) {
        fail(SVTK_ERR_CUDA, "no CUDA device: %s -- libsvtk_b200 has no CPU fallback", e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
        cudaGetLastError();
        return nullptr;
    }
    if (device < 0 || device >= n) { fail(SVTK_ERR_ARG, "device %d out of range [0,%d)", device, n); return nullptr; }
    if (cudaSetDevice(device) != cudaSuccess) { fail(SVTK_ERR_CUDA, "cudaSetDevice(%d) failed", device); return nullptr; }
    Ctx *c = new Ctx();
    c->device = device;
    const svtk_params def = {20000, 10000, 2000, 500, 5, 3};      // params.h:27-32
    c->params = params ? *params : def;
    if (cuda_stream) { c->stream = (cudaStream_t)cuda_stream; c->own_stream = false; }
    else {
        if (cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess) {
            fail(SVTK_ERR_CUDA, "cudaStreamCreate failed"); delete c; return nullptr;
        }
        c->own_stream = true;
    }
    return reinterpret_cast<svtk_ctx *>(c);
}

void svtk_destroy(svtk_ctx *ctx) {
    Ctx *c = reinterpret_cast<Ctx *>(ctx);
    if (!c) return;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    for (DevBuf *b : {&c->r_tid_off, &c->r_pos, &c->r_end, &c->r_ncig, &c->r_off, &c->r_cigar, &c->range, &c->ov_items,
                      &c->arena, &c->counters, &c->scratch4, &c->idx_events, &c->idx_chunks, &c->idx_segs, &c->idx_rend, &c->idx_ctrl, &c->pos_tiles, &c->pos_tile_off, &c->win_blob, &c->result, &c->tmp_a, &c->tmp_b, &c->tmp_c,
                      &c->poa_in, &c->poa_out, &c->poa_scratch, &c->poa_ctl, &c->bam_comp, &c->bam_infl, &c->bam_blocks,
                      &c->bam_splits, &c->bam_seg, &c->bam_rec, &c->bam_ctl, &c->bam_tok, &c->bam_ntok})
        b->release();
    for (PinBuf *b : {&c->stage[0], &c->stage[1], &c->h_idx, &c->h_win, &c->h_result, &c->poa_h, &c->h_audit}) b->release();
    for (int i = 0; i < 3; ++i) if (c->poa_ev[i]) cudaEventDestroy(c->poa_ev[i]);
    for (int i = 0; i < 5; ++i) if (c->bam_ev[i]) cudaEventDestroy(c->bam_ev[i]);
    c->bam_h.release(); c->bam_tab.release();
    for (int i = 0; i < 3; ++i) { c->bam_stage[i].release(); if (c->bam_stage_ev[i]) cudaEventDestroy(c->bam_stage_ev[i]); }
    for (int i = 0; i < 4; ++i) { if (c->bam_side_ev[i]) cudaEventDestroy(c->bam_side_ev[i]); if (c->bam_side[i]) cudaStreamDestroy(c->bam_side[i]); }
    for (int i = 0; i < 8; ++i) if (c->bam_chunk_ev[i]) cudaEventDestroy(c->bam_chunk_ev[i]);
    if (c->poa_side) cudaStreamDestroy(c->poa_side);
    for (int i = 0; i < 2; ++i) if (c->stage_ev[i]) cudaEventDestroy(c->stage_ev[i]);
    for (int i = 0; i < 4; ++i) if (c->prof_ev[i]) cudaEventDestroy(c->prof_ev[i]);
    if (c->own_stream) cudaStreamDestroy(c->stream);
    delete c;
}

int svtk_set_params(svtk_ctx *ctx, const svtk_params *p) {
    if (!ctx || !p) return fail(SVTK_ERR_ARG, "null argument");
    reinterpret_cast<Ctx *>(ctx)->params = *p;
    return 0;
}
int svtk_get_params(const svtk_ctx *ctx, svtk_params *p) {
    if (!ctx || !p) return fail(SVTK_ERR_ARG, "null argument");
    *p = reinterpret_cast<const Ctx *>(ctx)->params;
    return 0;
}

int svtk_load_reads(svtk_ctx *ctx, const svtk_reads *h) {
    Ctx *c = reinterpret_cast<Ctx *>(ctx);
    if (!c) return fail(SVTK_ERR_ARG, "null ctx");
    if (int rc = check_reads(h)) return rc;
    if (int rc = set_device(c)) return rc;
    c->have_reads = false;
    const size_t nr = (size_t)h->n_reads, nw = (size_t)h->n_words;
    if (int rc = c->r_tid_off.reserve(sizeof(long long) * (size_t)(h->n_tid + 1))) return rc;
    if (int rc = c->r_pos.reserve(4 * nr)) return rc;
    if (int rc = c->r_end.reserve(4 * nr)) return rc;
    if (int rc = c->r_ncig.reserve(4 * nr)) return rc;
    if (int rc = c->r_off.reserve(8 * nr)) return rc;
    if (int rc = c->r_cigar.reserve(4 * nw + 16)) return rc;   // walkers read whole uint4s: keep the tail readable
    c->h_tid_off.assign(h->n_tid + 1, 0);
    if (h->n_tid > 0) memcpy(c->h_tid_off.data(), h->tid_off, sizeof(int64_t) * (size_t)(h->n_tid + 1));
    if (c->h_tid_off.back() != h->n_reads) return fail(SVTK_ERR_ARG, "tid_off[n_tid] != n_reads");
    if (int rc = upload(c, c->r_tid_off.p, c->h_tid_off.data(), sizeof(long long) * (size_t)(h->n_tid + 1))) return rc;
    if (int rc = upload(c, c->r_pos.p, h->pos, 4 * nr)) return rc;
    if (int rc = upload(c, c->r_end.p, h->end, 4 * nr)) return rc;
    if (int rc = upload(c, c->r_ncig.p, h->n_cigar, 4 * nr)) return rc;
    if (int rc = upload(c, c->r_off.p, h->cigar_off, 8 * nr)) return rc;
    if (int rc = upload(c, c->r_cigar.p, h->cigar, 4 * nw)) return rc;
    c->reads = LaunchReads{h->n_tid, h->n_reads, (const long long *)c->r_tid_off.p, (const uint32_t *)c->r_pos.p,
                           (const uint32_t *)c->r_end.p, (const uint32_t *)c->r_ncig.p, (const unsigned long long *)c->r_off.p,
                           (const uint32_t *)c->r_cigar.p, 1, nullptr, nullptr};
    c->h2d_bytes += sizeof(long long) * (size_t)(h->n_tid + 1) + 20 * nr + 4 * nw;
    c->n_words = h->n_words;
    return finish_reads(c);
}

int svtk_attach_reads_device(svtk_ctx *ctx, const svtk_reads *d) {
    Ctx *c = reinterpret_cast<Ctx *>(ctx);
    if (!c) return fail(SVTK_ERR_ARG, "null ctx");
    return adopt_device_reads(c, d);
}

static LaunchWindows as_launch(const svtk_windows *w) {
    return LaunchWindows{w->n_win, w->chrom, w->start, w->end, w->pos, w->kind};
}

int svtk_refine_windows_device(svtk_ctx *ctx, const svtk_windows *w, int32_t *dev_result) {
    Ctx *c = reinterpret_cast<Ctx *>(ctx);
    if (!c || !w || (w->n_win > 0 && !dev_result)) return fail(SVTK_ERR_ARG, "null argument");
    if (c->pending) if (int rc = svtk_sync(ctx)) return rc;
    return launch_batch(c, as_launch(w), dev_result, MODE_INDEXED);
}

// ---- svtk_audit_device: one whole pass of the path (index build + every window), stream-ordered, no host wait.
// Each step copies its control block (index counters, error bits, overflow count) to its own pinned slot; svtk_sync reads the
// slots in order and redoes, through the synchronous path, any step whose index ran out of room or whose windows overflowed
// the register sort.  Up to AUDIT_SLOTS steps may be in flight; one more waits for the stream.
static const int AUDIT_SLOTS = 256;
static size_t audit_slot_bytes() { return (ctl_index_bytes() + 64 + 63) & ~(size_t)63; }

static int resolve_audit_steps(Ctx *c) {
    // the stream has drained.  Steps are independent passes over the same resident reads: redo the ones that need it.
    std::vector<Ctx::AuditStep> q;
    q.swap(c->audit_q);
    c->pending = false;
    c->index_fresh = false;
    const int st = index_ctrl_stride(), np = index_pools();
    for (size_t i = 0; i < q.size(); ++i) {
        const unsigned int *ctrl = (const unsigned int *)((const char *)c->h_audit.p + (size_t)q[i].slot * audit_slot_bytes());
        const unsigned int err = ctrl[(np + 1) * st];
        const int n_ov = *(const int *)((const char *)ctrl + ctl_index_bytes());
        if (i + 1 == q.size()) {                             // totals of the newest index (svtk_measure_index reads them)
            uint64_t ev = 0;
            for (int p = 0; p < np; ++p) ev += ctrl[p * st];
            if (index_fixed_slots()) for (int p = 0; p < np; ++p) ev += ctrl[p * st + 1];
            c->idx_n_events = ev; c->idx_n_segs = ctrl[np * st];
        }
        if (!err && n_ov <= 0) continue;
        c->index_valid = false;                              // rebuild (re-dimensioned if needed) and redo this step's windows
        if (int rc = launch_batch(c, q[i].win, q[i].result, MODE_INDEXED)) return rc;
        CU(cudaStreamSynchronize(c->stream));
        if (int rc = resolve_overflow(c)) return rc;
    }
    return 0;
}

int svtk_audit_device(svtk_ctx *ctx, const svtk_windows *w, int32_t *dev_result) {
    Ctx *c = reinterpret_cast<Ctx *>(ctx);
    if (!c || !w || (w->n_win > 0 && !dev_result)) return fail(SVTK_ERR_ARG, "null argument");
    if (!c->have_reads) return fail(SVTK_ERR_STATE, "no reads loaded");
    if (int rc = set_device(c)) return rc;
    if ((c->pending && c->audit_q.empty()) || (int)c->audit_q.size() >= AUDIT_SLOTS) if (int rc = svtk_sync(ctx)) return rc;
    if (int rc = c->h_audit.reserve(audit_slot_bytes() * AUDIT_SLOTS)) return rc;
    if (w->n_win <= 0) return build_index(c);
    const int slot = (int)c->audit_q.size();
    c->index_valid = false;                                  // a step always contains the whole job: rebuild the index
    if (int rc = launch_batch(c, as_launch(w), dev_result, MODE_INDEXED, (char *)c->h_audit.p + (size_t)slot * audit_slot_bytes())) return rc;
    c->audit_q.push_back(Ctx::AuditStep{as_launch(w), dev_result, slot});
    return 0;
}

int svtk_sync(svtk_ctx *ctx) {
    Ctx *c = reinterpret_cast<Ctx *>(ctx);
    if (!c) return fail(SVTK_ERR_ARG, "null ctx");
    if (int rc = set_device(c)) return rc;
    CU(cudaStreamSynchronize(c->stream));
    if (c->profile) if (int rc = collect_timing(c)) return rc;
    if (!c->audit_q.empty()) return resolve_audit_steps(c);
    return resolve_overflow(c);
}

uint64_t svtk_launch_count(const svtk_ctx *ctx) { return ctx ? reinterpret_cast<const Ctx *>(ctx)->launches : 0; }

int svtk_set_profiling(svtk_ctx *ctx, int on) {
    Ctx *c = reinterpret_cast<Ctx *>(ctx);
    if (!c) return fail(SVTK_ERR_ARG, "null ctx");
    if (int rc = set_device(c)) return rc;
    if (on && !c->prof_ev[0])
        for (int i = 0; i < 4; ++i) CU(cudaEventCreate(&c->prof_ev[i]));
    if (int rc = collect_timing(c)) return rc;
    c->profile = on != 0;
    c->prof_stride = on > 1 ? on : 1;
    c->prof_tick_idx = c->prof_tick_ref = 0;
    c->refine_ms = c->index_ms = 0.0; c->refine_launches = c->index_launches = 0;
    return 0;
}

int svtk_get_timing(svtk_ctx *ctx, svtk_timing *t) {
    Ctx *c = reinterpret_cast<Ctx *>(ctx);
    if (!c || !t) return fail(SVTK_ERR_ARG, "null argument");
    if (int rc = set_device(c)) return rc;
    if (int rc = collect_timing(c)) return rc;
    t->refine_kernel_ms = c->refine_ms;
    t->refine_kernel_launches = c->refine_launches;
    t->index_kernel_ms = c->index_ms;
    t->index_kernel_launches = c->index_launches;
    t->h2d_bytes = c->h2d_bytes;
    t->d2h_bytes = c->d2h_bytes;
    return 0;
}

// (re)build the signature index of the resident reads now, asynchronously on the ctx stream
int svtk_index_reads(svtk_ctx *ctx) {
    Ctx *c = reinterpret_cast<Ctx *>(ctx);
    if (!c) return fail(SVTK_ERR_ARG, "null ctx");
    if (!c->have_reads) return fail(SVTK_ERR_STATE, "no reads loaded");
    if (int rc = set_device(c)) return rc;
    if (c->pending) if (int rc = svtk_sync(ctx)) return rc;
    return build_index(c);
}

int svtk_measure_index(svtk_ctx *ctx, const svtk_windows *w, int32_t *dev_result, svtk_index_stats *st) {
    Ctx *c = reinterpret_cast<Ctx *>(ctx);
    if (!c || !w || !st) return fail(SVTK_ERR_ARG, "null argument");
    if (c->pending) if (int rc = svtk_sync(ctx)) return rc;
    c->index_valid = false;
    if (int rc = launch_batch(c, as_launch(w), dev_result, MODE_INDEXED_STATS)) return rc;
    if (int rc = svtk_sync(ctx)) return rc;
    unsigned long long cnt[5] = {0, 0, 0, 0, 0};
    if (w->n_win > 0) CU(cudaMemcpy(cnt, c->counters.p, sizeof cnt, cudaMemcpyDeviceToHost));
    const uint64_t nr = (uint64_t)c->reads.n_reads, nw = (uint64_t)c->n_words, nc = (uint64_t)c->idx.n_chunks;
    st->n_reads = nr; st->n_words = nw; st->n_chunks = nc;
    st->n_events = c->idx_n_events; st->n_segments = c->idx_n_segs;
    st->index_bytes = 4 * nw + 20 * nr + 16 * c->idx_n_events + 4 * nr + 16 * nc;
    st->n_windows = (uint64_t)w->n_win;
    st->refine_events_scanned = cnt[0]; st->refine_segments = cnt[1]; st->refine_reads_in_range = cnt[2];
    st->n_signatures = cnt[3]; st->n_overflow = cnt[4];
    st->refine_bytes = 16 * cnt[0] + 16 * cnt[1] + 28ull * (uint64_t)w->n_win;
    return 0;
}

int svtk_measure_windows(svtk_ctx *ctx, const svtk_windows *w, int32_t *dev_result, svtk_stats *st) {
    Ctx *c = reinterpret_cast<Ctx *>(ctx);
    if (!c || !w || !st) return fail(SVTK_ERR_ARG, "null argument");
    if (c->pending) if (int rc = svtk_sync(ctx)) return rc;
    if (int rc = launch_batch(c, as_launch(w), dev_result, MODE_WALKER_STATS)) return rc;
    if (int rc = svtk_sync(ctx)) return rc;
    unsigned long long cnt[4] = {0, 0, 0, 0};
    if (w->n_win > 0) CU(cudaMemcpy(cnt, c->counters.p, sizeof cnt, cudaMemcpyDeviceToHost));
    st->n_windows = (uint64_t)w->n_win;
    st->n_pairs = cnt[0]; st->n_words_visited = cnt[1]; st->n_signatures = cnt[2]; st->n_overflow = cnt[3];
    st->algorithmic_bytes = 4 * cnt[1] + 16 * cnt[0] + 4 * cnt[2] + 28ull * (uint64_t)w->n_win;
    return 0;
}

// host windows -> pinned blob -> one H2D; layout [chrom | start | end | pos | kind]
int svtk_refine_windows(svtk_ctx *ctx, const svtk_windows *w, int32_t *host_result) {
    Ctx *c = reinterpret_cast<Ctx *>(ctx);
    if (!c || !w) return fail(SVTK_ERR_ARG, "null argument");
    if (w->n_win < 0) return fail(SVTK_ERR_ARG, "n_win < 0");
    if (w->n_win == 0) return 0;
    if (!w->chrom || !w->start || !w->end || !w->pos || !w->kind || !host_result) return fail(SVTK_ERR_ARG, "null window array");
    if (!c->have_reads) return fail(SVTK_ERR_STATE, "no reads loaded (svtk_load_reads / svtk_attach_reads_device)");
    if (int rc = set_device(c)) return rc;
    if (c->pending) if (int rc = svtk_sync(ctx)) return rc;
    const size_t n = (size_t)w->n_win, n4 = (n * 4 + 15) & ~(size_t)15;
    const size_t blob = 4 * n4 + ((n + 15) & ~(size_t)15);
    if (int rc = c->h_win.reserve(blob)) return rc;
    if (int rc = c->win_blob.reserve(blob)) return rc;
    if (int rc = c->result.reserve(4 * n)) return rc;
    if (int rc = c->h_result.reserve(4 * n)) return rc;
    char *hb = (char *)c->h_win.p;
    memcpy(hb, w->chrom, 4 * n); memcpy(hb + n4, w->start, 4 * n); memcpy(hb + 2 * n4, w->end, 4 * n);
    memcpy(hb + 3 * n4, w->pos, 4 * n); memcpy(hb + 4 * n4, w->kind, n);
    CU(cudaMemcpyAsync(c->win_blob.p, hb, blob, cudaMemcpyHostToDevice, c->stream));
    char *db = (char *)c->win_blob.p;
    const LaunchWindows lw{w->n_win, (const int32_t *)db, (const uint32_t *)(db + n4), (const uint32_t *)(db + 2 * n4),
                           (const uint32_t *)(db + 3 * n4), (const uint8_t *)(db + 4 * n4)};
    if (int rc = launch_batch(c, lw, (int32_t *)c->result.p, MODE_INDEXED)) return rc;
    CU(cudaMemcpyAsync(c->h_result.p, c->result.p, 4 * n, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    if (*host_flags(c) > 0 || (c->index_fresh && ((const unsigned int *)c->h_idx.p)[(index_pools() + 1) * index_ctrl_stride()])) {
        // some windows left the fast path, or the fresh index ran out of room: finish them, then fetch again
        if (int rc = resolve_overflow(c)) return rc;
        CU(cudaMemcpyAsync(c->h_result.p, c->result.p, 4 * n, cudaMemcpyDeviceToHost, c->stream));
        CU(cudaStreamSynchronize(c->stream));
    }
    if (c->index_fresh) { c->index_fresh = false; unsigned int mp, er; index_totals(c, &c->idx_n_events, &c->idx_n_segs, &mp, &er); }
    c->pending = false;
    memcpy(host_result, c->h_result.p, 4 * n);
    c->h2d_bytes += blob;
    c->d2h_bytes += 4 * n + 4;
    return 0;
}

// ---- per-record seam (refinement.c:327-339)
int svtk_deletion(svtk_ctx *ctx, int chrom, svtk_interval begin, svtk_interval end, svtk_interval sv_inter, svtk_interval *res) {
    if (!res) return fail(SVTK_ERR_ARG, "null result");
    const int32_t ch[2] = {chrom, chrom};
    const uint32_t st[2] = {begin.start, end.start}, en[2] = {begin.end, end.end}, ps[2] = {sv_inter.start, sv_inter.end};
    const uint8_t kd[2] = {SVTK_DEL_START, SVTK_DEL_END};
    const svtk_windows w = {2, ch, st, en, ps, kd};
    int32_t r[2];
    if (int rc = svtk_refine_windows(ctx, &w, r)) return rc;
    res->start = (uint32_t)r[0]; res->end = (uint32_t)r[1];
    return 0;
}

int svtk_insertion(svtk_ctx *ctx, int chrom, svtk_interval begin, uint32_t pos, uint32_t *res_start) {
    if (!res_start) return fail(SVTK_ERR_ARG, "null result");
    const int32_t ch = chrom;
    const uint8_t kd = SVTK_INS;
    const svtk_windows w = {1, &ch, &begin.start, &begin.end, &pos, &kd};
    int32_t r;
    if (int rc = svtk_refine_windows(ctx, &w, &r)) return rc;
    *res_start = (uint32_t)r;
    return 0;
}

int svtk_inversion(svtk_ctx *ctx, int chrom, svtk_interval begin, svtk_interval end, svtk_interval sv_inter, svtk_interval *res) {
    if (!res) return fail(SVTK_ERR_ARG, "null result");
    const int32_t ch[2] = {chrom, chrom};
    const uint32_t st[2] = {begin.start, end.start}, en[2] = {begin.end, end.end}, ps[2] = {sv_inter.start, sv_inter.end};
    const uint8_t kd[2] = {SVTK_POINT, SVTK_POINT};
    const svtk_windows w = {2, ch, st, en, ps, kd};
    int32_t r[2];
    if (int rc = svtk_refine_windows(ctx, &w, r)) return rc;
    res->start = (uint32_t)r[0]; res->end = (uint32_t)r[1];
    return 0;
}

// ---- parity aids
int svtk_consensus_pos_batch(svtk_ctx *ctx, const int32_t *values, const int64_t *offsets, int32_t n_seg, const int32_t *pos,
                             int32_t *result) {
    Ctx *c = reinterpret_cast<Ctx *>(ctx);
    if (!c || !offsets || !pos || !result || n_seg < 0) return fail(SVTK_ERR_ARG, "bad argument");
    if (n_seg == 0) return 0;
    if (int rc = set_device(c)) return rc;
    const size_t total = (size_t)offsets[n_seg];
    if (total && !values) return fail(SVTK_ERR_ARG, "values is NULL");
    const int cap = consensus_segment_smem_ints();
    std::vector<long long> soff((size_t)n_seg, 0);
    unsigned long long scratch_ints = 0;
    for (int s = 0; s < n_seg; ++s) {
        const long long n = offsets[s + 1] - offsets[s];
        if (n < 0 || n > (1ll << 30)) return fail(SVTK_ERR_ARG, "segment %d has bad length", s);
        soff[s] = (long long)scratch_ints;
        if (n > cap) { unsigned long long p = 1; while (p < (unsigned long long)n) p <<= 1; scratch_ints += p; }
    }
    DevBuf d_val, d_off, d_pos, d_soff, d_res;
    int rc = 0;
    auto done = [&](int code) { d_val.release(); d_off.release(); d_pos.release(); d_soff.release(); d_res.release(); return code; };
    if ((rc = d_val.reserve(4 * total + 16)) || (rc = d_off.reserve(8 * (size_t)(n_seg + 1))) || (rc = d_pos.reserve(4 * (size_t)n_seg)) ||
        (rc = d_soff.reserve(8 * (size_t)n_seg)) || (rc = d_res.reserve(4 * (size_t)n_seg)) || (rc = c->arena.reserve(scratch_ints * 4)))
        return done(rc);
    if ((rc = upload(c, d_val.p, values, 4 * total)) || (rc = upload(c, d_off.p, offsets, 8 * (size_t)(n_seg + 1))) ||
        (rc = upload(c, d_pos.p, pos, 4 * (size_t)n_seg)) || (rc = upload(c, d_soff.p, soff.data(), 8 * (size_t)n_seg)))
        return done(rc);
    cudaError_t e = launch_consensus_segments((const int32_t *)d_val.p, (const long long *)d_off.p, (const int32_t *)d_pos.p, n_seg,
                                              c->params.consensus_interval_range, c->params.consensus_interval,
                                              c->params.consensus_min_count, (int *)c->arena.p, (const long long *)d_soff.p,
                                              (int32_t *)d_res.p, c->stream);
    c->launches += 1;
    if (e == cudaSuccess) e = cudaMemcpyAsync(result, d_res.p, 4 * (size_t)n_seg, cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    if (e != cudaSuccess) return done(fail(SVTK_ERR_CUDA, "consensus_pos_batch: %s", cudaGetErrorString(e)));
    return done(0);
}

static int upload_windows(Ctx *c, const svtk_windows *w, LaunchWindows *lw) {
    const size_t n = (size_t)w->n_win, n4 = (n * 4 + 15) & ~(size_t)15;
    const size_t blob = 4 * n4 + ((n + 15) & ~(size_t)15);
    if (int rc = c->h_win.reserve(blob)) return rc;
    if (int rc = c->win_blob.reserve(blob)) return rc;
    char *hb = (char *)c->h_win.p;
    memcpy(hb, w->chrom, 4 * n); memcpy(hb + n4, w->start, 4 * n); memcpy(hb + 2 * n4, w->end, 4 * n);
    memcpy(hb + 3 * n4, w->pos, 4 * n); memcpy(hb + 4 * n4, w->kind, n);
    CU(cudaMemcpyAsync(c->win_blob.p, hb, blob, cudaMemcpyHostToDevice, c->stream));
    char *db = (char *)c->win_blob.p;
    *lw = LaunchWindows{w->n_win, (const int32_t *)db, (const uint32_t *)(db + n4), (const uint32_t *)(db + 2 * n4),
                        (const uint32_t *)(db + 3 * n4), (const uint8_t *)(db + 4 * n4)};
    return 0;
}

int svtk_window_signatures(svtk_ctx *ctx, int kind, int chrom, svtk_interval inter, int32_t *sig_out, int64_t cap, int64_t *n_sig) {
    Ctx *c = reinterpret_cast<Ctx *>(ctx);
    if (!c || !n_sig || (cap > 0 && !sig_out)) return fail(SVTK_ERR_ARG, "bad argument");
    if (!c->have_reads) return fail(SVTK_ERR_STATE, "no reads loaded");
    if (int rc = set_device(c)) return rc;
    if (c->pending) if (int rc = svtk_sync(ctx)) return rc;
    const int32_t ch = chrom; const uint32_t ps = 0; const uint8_t kd = (uint8_t)kind;
    const svtk_windows w = {1, &ch, &inter.start, &inter.end, &ps, &kd};
    LaunchWindows lw;
    if (int rc = upload_windows(c, &w, &lw)) return rc;
    if (int rc = ensure_batch(c, 1)) return rc;
    c->flags_clean = false;                                  // this entry point uses the batch flags as its own counter
    CU(launch_plan(c->reads, lw, (int2 *)c->range.p, c->stream));
    c->launches += 1;
    long long buf_ints = 1 << 16;
    for (int attempt = 0; attempt < 3; ++attempt) {
        if (int rc = c->tmp_a.reserve((size_t)buf_ints * 4 + 16)) return rc;
        CU(launch_window_signatures(c->reads, lw, (const int2 *)c->range.p, (int *)c->tmp_a.p, (int)buf_ints, ov_flags(c), c->stream));
        c->launches += 1;
        int n = 0;
        CU(cudaMemcpyAsync(&n, ov_flags(c), 4, cudaMemcpyDeviceToHost, c->stream));
        CU(cudaStreamSynchronize(c->stream));
        long long p = 1; while (p < n) p <<= 1;
        if (p <= buf_ints) {
            *n_sig = n;
            const size_t k = (size_t)std::min<long long>(n, cap);
            if (k) CU(cudaMemcpy(sig_out, c->tmp_a.p, 4 * k, cudaMemcpyDeviceToHost));
            return 0;
        }
        buf_ints = p;
    }
    return fail(SVTK_ERR_STATE, "window_signatures did not converge");
}

int svtk_sliding_window_ins(svtk_ctx *ctx, int chrom, svtk_interval inter, int window_size, int slide_size, int32_t *best,
                            char *lines_buf, size_t cap) {
    Ctx *c = reinterpret_cast<Ctx *>(ctx);
    if (!c || !best) return fail(SVTK_ERR_ARG, "bad argument");
    if (window_size <= 0 || slide_size <= 0) return fail(SVTK_ERR_ARG, "window_size and slide_size must be positive (the reference loops forever otherwise)");
    if (!c->have_reads) return fail(SVTK_ERR_STATE, "no reads loaded");
    if (int rc = set_device(c)) return rc;
    if (c->pending) if (int rc = svtk_sync(ctx)) return rc;
    if (lines_buf && cap) lines_buf[0] = 0;
    *best = -1;
    // tiles exactly as sliding_window.c:12-15 forms them (uint32 arithmetic)
    std::vector<uint32_t> st, en;
    for (uint32_t s = inter.start; s < inter.end;) {
        uint32_t e = s + (uint32_t)window_size;
        if (e > inter.end) e = inter.end;
        st.push_back(s); en.push_back(e);
        if (st.size() > (1u << 22)) return fail(SVTK_ERR_ARG, "more than 4M tiles");
        const uint32_t nx = s + (uint32_t)window_size;
        if (nx <= s) break;                      // uint32 wrap: the reference would never terminate
        s = nx;
    }
    const int nt = (int)st.size();
    if (nt == 0) return 0;
    std::vector<int32_t> ch((size_t)nt, chrom);
    std::vector<uint32_t> ps((size_t)nt, 0);
    std::vector<uint8_t> kd((size_t)nt, (uint8_t)SVTK_INS);
    const svtk_windows w = {nt, ch.data(), st.data(), en.data(), ps.data(), kd.data()};
    LaunchWindows lw;
    if (int rc = upload_windows(c, &w, &lw)) return rc;
    if (int rc = ensure_batch(c, nt)) return rc;
    c->flags_clean = false;                                  // this entry point uses the batch flags as its own counter
    if (int rc = c->tmp_b.reserve(sizeof(int2) * (size_t)nt)) return rc;
    CU(launch_plan(c->reads, lw, (int2 *)c->range.p, c->stream));
    c->launches += 1;
    int *need_d = ov_flags(c);
    long long stride = 0;
    std::vector<int2> out((size_t)nt);
    for (int attempt = 0; attempt < 2; ++attempt) {
        CU(cudaMemsetAsync(need_d, 0, 4, c->stream));
        int *gseg = nullptr;
        if (stride) {
            if (int rc = c->arena.reserve((size_t)stride * 4 * (size_t)nt)) return rc;
            gseg = (int *)c->arena.p;
        }
        CU(launch_slide_tiles(c->reads, lw, (const int2 *)c->range.p, window_size, slide_size, c->params.consensus_min_count, gseg, stride,
                              (int2 *)c->tmp_b.p, need_d, c->stream));
        c->launches += 1;
        int need = 0;
        CU(cudaMemcpyAsync(&need, need_d, 4, cudaMemcpyDeviceToHost, c->stream));
        CU(cudaMemcpyAsync(out.data(), c->tmp_b.p, sizeof(int2) * (size_t)nt, cudaMemcpyDeviceToHost, c->stream));
        CU(cudaStreamSynchronize(c->stream));
        if (need == 0) break;
        if (attempt == 1) return fail(SVTK_ERR_STATE, "sliding-window tile segment overflow");
        stride = 1; while (stride < need) stride <<= 1;
    }
    int best_all = -1, max_all = 0;
    std::string lines;
    char tmp[160];
    for (int i = 0; i < nt; ++i) {
        if (out[i].y > 0 && out[i].x != -1) {    // sliding_window.c:86-92
            snprintf(tmp, sizeof tmp, "INS Discovery in window [%d, %d] at position %d with support %d\n", (int)st[i], (int)en[i], out[i].x, out[i].y);
            lines += tmp;
            if (out[i].y > max_all) { max_all = out[i].y; best_all = out[i].x; }
        }
    }
    *best = best_all;
    if (lines_buf && cap) { const size_t k = std::min(cap - 1, lines.size()); memcpy(lines_buf, lines.data(), k); lines_buf[k] = 0; }
    return 0;
}

}  // extern "C"
