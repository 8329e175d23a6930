// svtk_index.cu -- the streaming form of the audt refinement path (sm_100a).
//
// The reference walks a read's CIGAR once per window that sees it (refinement.c:117-160 inside every refine_*
// call), i.e. every CIGAR word is decoded ~7 times at 30x.  What a window can take from a read depends on the
// window only through two comparisons:
//     - a D/I op is emitted iff the reference cursor BEFORE it has not passed W.end   (stop rule, :141-144)
//     - the clip rules compare the read's start / final cursor with [W.start, W.end]  (:147-159, :210-221)
// so the walk itself is window independent.  This file splits the path accordingly:
//
//   k_index_reads_defer  ONE coalesced pass over all CIGAR words (the product kernel): a warp owns a chunk of 8 consecutive
//                     reads and streams their packed CIGAR words, 256 per step, through a cp.async ring; a step is decode +
//                     warp scan and leaves only per-lane marks; one resolve pass per 8 steps turns the marks into the
//                     chunk's signature events {cursor, len|type, read, endpos} -- D > 50, I >= 50, trailing / leading soft
//                     clip -- plus every read's final cursor.  No block-level barrier anywhere.
//                     (two superseded forms of this pass and their measurements: tools/ab/index_variants.cuh)
//   k_refine_warp     one WARP per window: 32-ary search of the read range the BAM iterator would yield
//                     (hts.c:3049-3090, 3905-3921), gather of the covering chunks' events under the two rules
//                     above, register bitonic sort (<= 256 signatures), consensus_pos (refinement.c:41-101).
//   k_refine_block    windows with more signatures: one CTA each, shared-memory or global segment.
//
// Results are bit-identical to the per-(window, read) walker in svtk_kernels.cu (kept as the instrumented
// cross-check behind svtk_measure_windows).
#include "svtk_device.cuh"
#include "svtk_launch.h"


namespace svtk {

constexpr int EV_DEL = 0, EV_INS = 1, EV_TCLIP = 2, EV_LCLIP = 3;

#ifndef SVTK_IDX_WARPS
#define SVTK_IDX_WARPS 4
#endif
constexpr int IDX_WARPS = SVTK_IDX_WARPS;    // warps (= chunks) per CTA of k_index_reads
#ifndef SVTK_IDX_WPT
#define SVTK_IDX_WPT 8
#endif
constexpr int IDX_WPT = SVTK_IDX_WPT;        // CIGAR words per lane and step
constexpr int IDX_TILE = 32 * IDX_WPT;      // 256 words = 1 KB per warp step
constexpr int CTRL_STRIDE = 32;             // u32 words between counters: one 128-byte line each
constexpr int CTRL_SEG = IDX_POOLS * CTRL_STRIDE, CTRL_ERR = CTRL_SEG + CTRL_STRIDE;

struct DevIndex {
    uint4 *events;       // {x, len << 4 | type, read index, read endpos}; IDX_POOLS sub-pools of pool_cap events
    uint4 *chunks;       // per chunk of IDX_READS reads: newest segment {base, count, next segment or -1, first pos}
    uint4 *segs;         // older segments of chunks that flushed more than once
    uint32_t *rend;      // per read: reference cursor after the whole CIGAR
    unsigned int *ctrl;  // [p * CTRL_STRIDE] top of sub-pool p, [CTRL_SEG] segment top, [CTRL_ERR] error bits (1 events, 2 segments)
    unsigned int pool_cap, seg_cap;
    int n_chunks;
    unsigned int *next_ctrl;   // the next build's control block: cleared here, off the host's critical path
    int next_words;
};

static DevIndex to_dev(const LaunchIndex &x) {
    DevIndex d;
    d.events = reinterpret_cast<uint4 *>(x.events); d.chunks = reinterpret_cast<uint4 *>(x.chunks);
    d.segs = reinterpret_cast<uint4 *>(x.segs); d.rend = x.rend; d.ctrl = x.ctrl; d.pool_cap = x.pool_cap; d.seg_cap = x.seg_cap;
    d.n_chunks = x.n_chunks; d.next_ctrl = x.next_ctrl; d.next_words = x.next_words;
    return d;
}
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

// ---------------------------------------------------------------------------------------------------------
// high half of a * b, plus c: one IMAD.HI on the FMA pipe (the decode loop is ALU-pipe bound otherwise)
__device__ __forceinline__ uint32_t mad_hi_u32(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t d;
    asm("mad.hi.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

// A segment keeps its events in two classes, one after the other: first the ones deletion windows look at (D > 50, trailing /
// leading soft clip), then the insertions.  A window reads only its own class -- half the events at an even DEL / INS mix.
// Descriptor .y = count of the first class | count of the second << 16 (a segment holds <= IDX_DEVCAP events).
__device__ __forceinline__ uint32_t seg_count(uint32_t y) { return (y & 0xffffu) + (y >> 16); }
__device__ __forceinline__ void seg_class(uint32_t base, uint32_t y, bool ins, uint32_t &cbase, uint32_t &ccount) {
    cbase = base + (ins ? (y & 0xffffu) : 0u);
    ccount = ins ? (y >> 16) : (y & 0xffffu);
}

// s.ev[0 .. nev) -> X.events[fb ..], stably partitioned by class; returns the packed counts (uniform).  nev <= 64.
template <class SH>
__device__ __forceinline__ uint32_t index_store_partitioned(SH &s, const DevIndex &X, uint32_t fb, int nev) {
    const int lane = lane_id();
    const unsigned below = (1u << lane) - 1u;
    uint4 e0 = make_uint4(0u, 0u, 0u, 0u), e1 = e0;
    const bool h0 = lane < nev, h1 = lane + 32 < nev;
    if (h0) e0 = s.ev[lane];
    if (h1) e1 = s.ev[lane + 32];
    const bool b0 = h0 && (e0.y & 15u) == (uint32_t)EV_INS, b1 = h1 && (e1.y & 15u) == (uint32_t)EV_INS;
    const unsigned mb0 = __ballot_sync(FULL, b0), mb1 = __ballot_sync(FULL, b1);
    const unsigned ma0 = __ballot_sync(FULL, h0 && !b0), ma1 = __ballot_sync(FULL, h1 && !b1);
    const uint32_t nA = (uint32_t)(__popc(ma0) + __popc(ma1)), nB = (uint32_t)(__popc(mb0) + __popc(mb1));
    if (h0) X.events[(size_t)fb + (b0 ? nA + (uint32_t)__popc(mb0 & below) : (uint32_t)__popc(ma0 & below))] = e0;
    if (h1) X.events[(size_t)fb + (b1 ? nA + (uint32_t)(__popc(mb0) + __popc(mb1 & below)) : (uint32_t)(__popc(ma0) + __popc(ma1 & below)))] = e1;
    return nA | nB << 16;
}

// append the warp's staged events to the chunk as one more segment; `head` (uniform) is the chunk's newest segment
template <class SH>
__device__ __forceinline__ void index_flush_warp(SH &s, const DevIndex &X, int pool, int nev, uint4 &head) {
    const int lane = lane_id();
    __syncwarp();
    uint32_t fb = 0xffffffffu;
    uint4 nh = head;
    if (lane == 0) {
        const unsigned int b = atomicAdd(&X.ctrl[pool * CTRL_STRIDE], (unsigned int)nev);
        if ((unsigned long long)b + (unsigned long long)nev > (unsigned long long)X.pool_cap) atomicOr(&X.ctrl[CTRL_ERR], 1u);
        else {
            uint32_t next = 0xffffffffu;
            bool ok = true;
            if (head.y != 0u) {                              // push the previous segment down the chain
                const unsigned int si = atomicAdd(&X.ctrl[CTRL_SEG], 1u);
                if (si >= X.seg_cap) { atomicOr(&X.ctrl[CTRL_ERR], 2u); ok = false; }
                else { X.segs[si] = head; next = si; }
            }
            if (ok) { fb = (uint32_t)pool * X.pool_cap + b; nh = make_uint4(fb, (uint32_t)nev, next, head.w); }
        }
    }
    fb = __shfl_sync(FULL, fb, 0);
    nh.x = __shfl_sync(FULL, nh.x, 0); nh.y = __shfl_sync(FULL, nh.y, 0); nh.z = __shfl_sync(FULL, nh.z, 0);
    if (fb != 0xffffffffu) nh.y = index_store_partitioned(s, X, fb, nev);
    head = nh;
    __syncwarp();
}

// ---------------------------------------------------------------------------------------------------------
// k_index_reads_defer: the same index with the per-step work cut to decode + scan.  A step leaves only two things behind,
// per lane: the cursor before its first word (lanebase) and the 8-bit mask of its signature words (candb).  Every IDX_NS
// steps (and at the end of the chunk) ONE resolve pass turns the marks into events, one item per lane: the item re-reads its
// lane's <= 7 preceding words (L1/L2 hits) to rebuild the cursor at the word.  The reads' first words are items of the same
// pass (they give base[r]).  Against k_index_reads this drops, from every step, the prefix array in shared memory, the second
// scan, the per-read base lookup and the divergent candidate loop.  Same events, same segments, same rend.
static_assert(IDX_WPT == 8, "the deferred marks are one byte per lane and step");
#ifndef SVTK_IDX_NS
#define SVTK_IDX_NS 8
#endif
constexpr int IDX_NS = SVTK_IDX_NS;          // steps between two resolve passes (8: 2 048 words); 4 or 8
static_assert(IDX_NS == 4 || IDX_NS == 8, "a lane owns IDX_NS mark bytes: one or two words");
constexpr int IDX_LIST = IDX_TILE + IDX_READS;
#ifndef SVTK_IDX_STAGE
#define SVTK_IDX_STAGE 2       // steps each lane keeps in flight through the cp.async ring (0: two-deep register pipeline).  Measured on B200,
                               // C2: 2 -> 35.8 us (10 CTAs/SM), 3 -> 39.0 (8), 4 -> 40.6 (7), 6 -> 46.8 (5): resident warps beat bytes in flight
#endif
constexpr int IDX_STAGE = SVTK_IDX_STAGE;
#ifndef SVTK_IDX_FIXK
#define SVTK_IDX_FIXK 32        // event slots every chunk owns behind the sub-pools (<= 32: one store per lane); 0: always allocate
#endif
constexpr int IDX_FIXK = SVTK_IDX_FIXK;
static_assert(IDX_FIXK <= 32, "one lane per fixed slot");
#ifndef SVTK_IDX_DEVCAP
#define SVTK_IDX_DEVCAP 64
#endif
constexpr int IDX_DEVCAP = SVTK_IDX_DEVCAP;  // events a warp stages between flushes (>= 32 + 2 * IDX_READS: one resolve round, the clip events)
static_assert(IDX_DEVCAP >= 32 + 2 * IDX_READS, "event staging too small");
static_assert(IDX_DEVCAP <= 64, "index_store_partitioned handles two events per lane");

struct __align__(16) IndexDeferShared {
    uint4 ev[IDX_DEVCAP];
    uint32_t lanebase[IDX_NS * 32];          // [step * 32 + lane]: cursor (chunk relative) before the lane's first word of the step
    uint32_t candw[IDX_NS * 8];              // bytes [step * 32 + lane]: which of the lane's 8 words are signatures
    uint32_t rel[IDX_READS + 4];
    uint32_t pos[IDX_READS], endp[IDX_READS];
    uint32_t base[IDX_READS + 1];
    uint16_t list[IDX_LIST];                 // items of one resolve pass: word id (11 bits) | read << 11 | 0x8000 for a read's first word
#if SVTK_IDX_STAGE
    __align__(16) uint4 ring[IDX_STAGE * (IDX_TILE / 4)];   // cp.async staging: IDX_STAGE steps of 1 KB, lane l owns bytes [32 l, 32 l + 32) of each
#endif
};

// inclusive warp scan on the shuffle's own in-range predicate: SHFL + predicated IADD per level
__device__ __forceinline__ uint32_t warp_inclusive_scan_p(uint32_t v) {
    asm("{\n\t.reg .pred p;\n\t.reg .u32 t;\n\t"
        "shfl.sync.up.b32 t|p, %0, 1, 0, 0xffffffff;\n\t@p add.u32 %0, %0, t;\n\t"
        "shfl.sync.up.b32 t|p, %0, 2, 0, 0xffffffff;\n\t@p add.u32 %0, %0, t;\n\t"
        "shfl.sync.up.b32 t|p, %0, 4, 0, 0xffffffff;\n\t@p add.u32 %0, %0, t;\n\t"
        "shfl.sync.up.b32 t|p, %0, 8, 0, 0xffffffff;\n\t@p add.u32 %0, %0, t;\n\t"
        "shfl.sync.up.b32 t|p, %0, 16, 0, 0xffffffff;\n\t@p add.u32 %0, %0, t;\n\t}"
        : "+r"(v));
    return v;
}

// {advance multiplier, signature threshold} of a CIGAR word's op: one LOP3 for the address, one LDS.64
__device__ __forceinline__ uint2 optab_lookup(uint32_t tab, uint32_t word) {
    uint2 ot;
    asm("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(ot.x), "=r"(ot.y) : "r"(((word << 3) & 0x78u) | tab));
    return ot;
}

// one resolve pass over the marks of buffered steps: all of them (st < 0; needs <= IDX_TILE marks) or step st alone
__device__ __forceinline__ void index_resolve(IndexDeferShared &s, const DevIndex &X, const uint32_t *__restrict__ cg, int pool, int &nev,
                                              uint4 &head, uint32_t tb0, int st, int nsteps, uint32_t head_pad, uint32_t span, int nr,
                                              long long r0, uint32_t tab) {
    const int lane = lane_id();
    const unsigned below = (1u << lane) - 1u;
    unsigned long long bits;
    uint32_t base_id, lo_w, hi_w;
    if (st < 0) {
        if constexpr (IDX_NS == 8) {
            const uint2 v = reinterpret_cast<const uint2 *>(s.candw)[lane];
            bits = (unsigned long long)v.x | ((unsigned long long)v.y << 32);
        } else bits = s.candw[lane];
        base_id = (uint32_t)lane * (8u * IDX_NS); lo_w = tb0; hi_w = tb0 + (uint32_t)nsteps * IDX_TILE;
    } else {
        bits = reinterpret_cast<const uint8_t *>(s.candw)[st * 32 + lane];
        base_id = (uint32_t)(st * 32 + lane) * 8u; lo_w = tb0 + (uint32_t)st * IDX_TILE; hi_w = lo_w + IDX_TILE;
    }
    const int cnt = __popcll(bits);
    uint32_t rel = 0xffffffffu;
    if (lane < nr) rel = s.rel[lane];
    const bool isstart = rel >= lo_w && rel < hi_w && rel < span;
    const unsigned ms = __ballot_sync(FULL, isstart);
    const int ns = __popc(ms);
    if (isstart) s.list[__popc(ms & below)] = (uint16_t)(0x8000u | ((uint32_t)lane << 11) | (rel - tb0));
    int incl = cnt;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const int o = __shfl_up_sync(FULL, incl, d); if (lane >= d) incl += o; }
    const int nc = __shfl_sync(FULL, incl, 31);
    int slot = ns + incl - cnt;
    while (bits) { const int b = __ffsll((long long)bits) - 1; bits &= bits - 1; s.list[slot++] = (uint16_t)(base_id + (uint32_t)b); }
    __syncwarp();
    const int total = ns + nc;
    for (int j0 = 0; j0 < total; j0 += 32) {
        const int j = j0 + lane;
        const int first = j0 ? j0 : ns;                      // first signature item of this round
        const int nround = min(total, j0 + 32) - first;
        if (nev + nround > IDX_DEVCAP) { index_flush_warp(s, X, pool, nev, head); nev = 0; }
        const bool act = j < total;
        uint32_t it = 0x8000u, c = 0, gi = 0;
        if (act) {
            it = s.list[j];
            const uint32_t id = it & 2047u, k = id & 7u;
            gi = tb0 + id;
            const uint32_t g0 = gi - k;                      // the owning lane's first word of that step (32-byte aligned)
            const uint4 *src = reinterpret_cast<const uint4 *>(cg + g0);
            uint4 a = make_uint4(0u, 0u, 0u, 0u), b = make_uint4(0u, 0u, 0u, 0u);
            if (k > 0) a = __ldg(src);
            if (k > 4) b = __ldg(src + 1);
            const uint32_t w[7] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z};
            c = s.lanebase[id >> 3];
#pragma unroll
            for (int q = 0; q < 7; ++q)
                if ((uint32_t)q < k && g0 + q >= head_pad) c = mad_hi_u32(w[q], optab_lookup(tab, w[q]).x, c);
            if (it & 0x8000u) s.base[(it >> 11) & 7u] = c;
        }
        __syncwarp();
        if (!(it & 0x8000u)) {
            const uint4 ra = *reinterpret_cast<const uint4 *>(&s.rel[0]), rb = *reinterpret_cast<const uint4 *>(&s.rel[4]);
            const int r = (int)(ra.y <= gi) + (int)(ra.z <= gi) + (int)(ra.w <= gi) + (int)(rb.x <= gi) + (int)(rb.y <= gi) +
                          (int)(rb.z <= gi) + (int)(rb.w <= gi);
            const uint32_t word = __ldg(cg + gi);
            const uint32_t x = s.pos[r] + (c - s.base[r]);
            const uint32_t type = ((word & 15u) == (uint32_t)OP_DEL) ? EV_DEL : EV_INS;
            s.ev[nev + j - first] = make_uint4(x, (word & ~15u) | type, (uint32_t)(r0 + r), s.endp[r]);
        }
        nev += nround;
    }
    __syncwarp();
}

__device__ __forceinline__ void index_resolve_all(IndexDeferShared &s, const DevIndex &X, const uint32_t *__restrict__ cg, int pool, int &nev,
                                                  uint4 &head, uint32_t tb0, int nsteps, uint32_t head_pad, uint32_t span, int nr,
                                                  long long r0, uint32_t tab) {
    const int lane = lane_id();
    __syncwarp();
    int mine;
    if constexpr (IDX_NS == 8) { const uint2 v = reinterpret_cast<const uint2 *>(s.candw)[lane]; mine = __popc(v.x) + __popc(v.y); }
    else mine = __popc(s.candw[lane]);
    const int tot = __reduce_add_sync(FULL, mine);
    if (tot <= IDX_TILE) index_resolve(s, X, cg, pool, nev, head, tb0, -1, nsteps, head_pad, span, nr, r0, tab);
    else
        for (int st = 0; st < nsteps; ++st) index_resolve(s, X, cg, pool, nev, head, tb0, st, nsteps, head_pad, span, nr, r0, tab);
    if constexpr (IDX_NS == 8) reinterpret_cast<uint2 *>(s.candw)[lane] = make_uint2(0u, 0u);
    else s.candw[lane] = 0u;
    __syncwarp();
}

#ifndef SVTK_IDX_DMINB
#define SVTK_IDX_DMINB 10       // 48 registers, 20 KB of shared memory per CTA: 40 warps per SM (12 CTAs at 40 registers spill: 43 us)
#endif
__global__ void __launch_bounds__(IDX_WARPS * 32, SVTK_IDX_DMINB) k_index_reads_defer(DevReads R, DevIndex X) {
    __shared__ IndexDeferShared sh[IDX_WARPS];
    __shared__ __align__(128) uint2 optab[16];               // one per CTA: every warp writes the same 16 entries, then reads after its own __syncwarp
    const int lane = lane_id(), warp = threadIdx.x >> 5;
    if (blockIdx.x == 0 && X.next_ctrl)                      // nobody touches that block before the next build (stream order)
        for (int i = threadIdx.x; i < X.next_words; i += IDX_WARPS * 32) X.next_ctrl[i] = 0u;
    const long long c = (long long)blockIdx.x * IDX_WARPS + warp;
    if (c >= X.n_chunks) return;
    IndexDeferShared &s = sh[warp];
    const int pool = (int)(blockIdx.x % IDX_POOLS);
    const long long r0 = c * IDX_READS;
    const int nr = (int)min((long long)IDX_READS, R.n_reads - r0);
    unsigned long long off = 0;
    if (lane <= nr) {
        const long long r = r0 + lane;
        off = (lane < nr) ? __ldg(R.cigar_off + r) : __ldg(R.cigar_off + r - 1) + __ldg(R.n_cigar + r - 1);
        if (lane < nr) { s.pos[lane] = __ldg(R.pos + r); s.endp[lane] = __ldg(R.end + r); }
    }
    const unsigned long long w0 = __shfl_sync(FULL, off, 0);
    const unsigned long long w0a = w0 & ~3ull;
    if (lane < IDX_READS + 4) s.rel[lane] = lane <= nr ? (uint32_t)(off - w0a) : 0xffffffffu;
    if (lane < 16) {
        const uint32_t adv = (lane == OP_INS || lane == OP_SOFT) ? 0u : (1u << 28);
        const uint32_t thr = lane == OP_INS ? ((uint32_t)SV_MIN_LENGTH << 4 | OP_INS)
                           : lane == OP_DEL ? ((uint32_t)(SV_MIN_LENGTH + 1) << 4 | OP_DEL) : 0xffffffffu;
        optab[lane] = make_uint2(adv, thr);
    }
    if constexpr (IDX_NS == 8) reinterpret_cast<uint2 *>(s.candw)[lane] = make_uint2(0u, 0u);
    else s.candw[lane] = 0u;
    __syncwarp();
    const uint32_t head_pad = (uint32_t)(w0 - w0a);
    const uint32_t span = s.rel[nr];
    const uint32_t *__restrict__ cg = R.cigar + w0a;
    uint4 head = make_uint4(0u, 0u, 0xffffffffu, s.pos[0]);
    uint8_t *candb = reinterpret_cast<uint8_t *>(s.candw);
    const uint32_t tab = (uint32_t)__cvta_generic_to_shared(optab);

    uint32_t run = 0;       // cursor advance of all words before the current step (uniform)
    int nev = 0;            // staged events (uniform)
    int sidx = 0;           // buffered steps (uniform)
#if SVTK_IDX_STAGE
    // Asynchronous staging: every lane keeps IDX_STAGE steps of its own 32-byte pieces in flight with 16-byte cp.async copies
    // into a per-warp ring in shared memory (no registers held by loads in flight: the kernel is bound by bytes in flight, not
    // by issue slots).  A lane reads back only what it copied itself, so cp.async.wait_group is the only synchronisation.  A
    // piece that starts at or beyond `span` is not copied; the stale ring contents are blanked by the edge rule below.
    const uint32_t ring = (uint32_t)__cvta_generic_to_shared(s.ring) + (uint32_t)lane * 32u;
    const uint32_t *__restrict__ np = cg + lane * IDX_WPT;
    uint32_t ni = (uint32_t)lane * IDX_WPT;                  // word index of the lane's next piece to copy
    int wslot = 0;
    auto issue = [&]() {
        const uint32_t dst = ring + (uint32_t)wslot * (IDX_TILE * 4u);
        if (ni < span) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(np) : "memory");
        if (ni + 4u < span) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + 16u), "l"(np + 4) : "memory");
        asm volatile("cp.async.commit_group;" ::: "memory");
        np += IDX_TILE; ni += IDX_TILE;
        wslot = (wslot + 1 == IDX_STAGE) ? 0 : wslot + 1;
    };
#pragma unroll
    for (int t = 0; t < IDX_STAGE; ++t) issue();
    uint32_t i0 = (uint32_t)lane * IDX_WPT;
    for (uint32_t tb = 0; tb < span; tb += IDX_TILE, i0 += IDX_TILE) {
        asm volatile("cp.async.wait_group %0;" ::"n"(IDX_STAGE - 1) : "memory");
        uint32_t w[IDX_WPT];
        {
            const uint32_t src = ring + (uint32_t)wslot * (IDX_TILE * 4u);   // the oldest slot == the one the next copy reuses
            asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]) : "r"(src) : "memory");
            asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(w[4]), "=r"(w[5]), "=r"(w[6]), "=r"(w[7]) : "r"(src + 16u) : "memory");
        }
        issue();
        // edge steps: blank the neighbouring chunks' words ("0M", inert).  head_pad <= 3: only lane 0's first words of step 0
        if (i0 == 0u) { if (head_pad > 0u) w[0] = 0u; if (head_pad > 1u) w[1] = 0u; if (head_pad > 2u) w[2] = 0u; }
        if (tb + IDX_TILE > span) {
            const int rem = (int)(span - i0);                // < 8: words from `rem` on are not this chunk's (<= 0: all of them)
#pragma unroll
            for (int k = 0; k < IDX_WPT; ++k) if (k >= rem) w[k] = 0u;
        }
        uint32_t sum = 0;
        unsigned candm = 0;
#pragma unroll
        for (int k = 0; k < IDX_WPT; ++k) {
            const uint2 ot = optab_lookup(tab, w[k]);
            sum = mad_hi_u32(w[k], ot.x, sum);               // += len for ops that advance (refinement.c:137)
            if (w[k] >= ot.y) candm |= 1u << k;              // I >= 50 (refinement.c:299), D > 50 (refinement.c:124)
        }
        const uint32_t incl = warp_inclusive_scan_p(sum);
        s.lanebase[sidx * 32 + lane] = run + (incl - sum);
        candb[sidx * 32 + lane] = (uint8_t)candm;
        run += __shfl_sync(FULL, incl, 31);
        if (++sidx == IDX_NS) {
            index_resolve_all(s, X, cg, pool, nev, head, tb + IDX_TILE - IDX_NS * IDX_TILE, IDX_NS, head_pad, span, nr, r0, tab);
            sidx = 0;
        }
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
#else
    // software pipeline: the next step's 32 bytes per lane are in flight while this step is decoded.  A piece that starts at
    // or beyond `span` is not loaded; its stale register contents are blanked by the edge rule below (every word >= span).
    uint4 A0 = make_uint4(0u, 0u, 0u, 0u), A1 = A0, B0 = A0, B1 = A0;
    const uint4 *__restrict__ np = reinterpret_cast<const uint4 *>(cg + lane * IDX_WPT);
    uint32_t ni = (uint32_t)lane * IDX_WPT;                  // word index of the lane's next piece
    if (ni < span) A0 = __ldg(np);
    if (ni + 4u < span) A1 = __ldg(np + 1);
    // one step: decode the pieces (c0, c1) while the next step's pieces load into (n0, n1)
    auto step = [&](const uint4 &c0, const uint4 &c1, uint4 &n0, uint4 &n1, const uint32_t tb) {
        const uint32_t i0 = ni;
        uint32_t w[IDX_WPT] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
        np += IDX_TILE / 4; ni += IDX_TILE;
        if (ni < span) n0 = __ldg(np);
        if (ni + 4u < span) n1 = __ldg(np + 1);
        if (tb == 0 || tb + IDX_TILE > span) {               // edge steps: blank the neighbouring chunks' words ("0M", inert)
#pragma unroll
            for (int k = 0; k < IDX_WPT; ++k) { const uint32_t gi = i0 + k; if (gi < head_pad || gi >= span) w[k] = 0u; }
        }
        uint32_t sum = 0;
        unsigned candm = 0;
#pragma unroll
        for (int k = 0; k < IDX_WPT; ++k) {
            const uint2 ot = optab_lookup(tab, w[k]);
            sum = mad_hi_u32(w[k], ot.x, sum);               // += len for ops that advance (refinement.c:137)
            if (w[k] >= ot.y) candm |= 1u << k;              // I >= 50 (refinement.c:299), D > 50 (refinement.c:124)
        }
        const uint32_t incl = warp_inclusive_scan_p(sum);
        s.lanebase[sidx * 32 + lane] = run + (incl - sum);
        candb[sidx * 32 + lane] = (uint8_t)candm;
        run += __shfl_sync(FULL, incl, 31);
        ++sidx;
    };
    // two steps per trip (the piece registers ping-pong, no copies); IDX_NS is even, so a full buffer is seen after the pair
    for (uint32_t tb = 0; tb < span; tb += 2 * IDX_TILE) {
        step(A0, A1, B0, B1, tb);
        if (tb + IDX_TILE < span) step(B0, B1, A0, A1, tb + IDX_TILE);
        if (sidx == IDX_NS) {
            index_resolve_all(s, X, cg, pool, nev, head, tb + 2 * IDX_TILE - IDX_NS * IDX_TILE, IDX_NS, head_pad, span, nr, r0, tab);
            sidx = 0;
        }
    }
#endif  // SVTK_IDX_STAGE
    if (sidx) index_resolve_all(s, X, cg, pool, nev, head, (span + IDX_TILE - 1) / IDX_TILE * IDX_TILE - (uint32_t)sidx * IDX_TILE, sidx,
                                head_pad, span, nr, r0, tab);
    // reads that start at the very end of the chunk's span (empty CIGARs) and the end-of-span sentinel
    if (lane <= nr && s.rel[lane] >= span) s.base[lane] = run;
    __syncwarp();
    if (nev + 2 * IDX_READS > IDX_DEVCAP) { index_flush_warp(s, X, pool, nev, head); nev = 0; }
    bool tclip = false, lclip = false;
    uint32_t rendv = 0;
    if (lane < nr) {
        const uint32_t n = s.rel[lane + 1] - s.rel[lane];
        rendv = s.pos[lane];
        if (n > 0) {
            rendv += s.base[lane + 1] - s.base[lane];
            const uint32_t first = __ldg(cg + s.rel[lane]), last = __ldg(cg + s.rel[lane + 1] - 1);
            tclip = (last & 15u) == (uint32_t)OP_SOFT;       // refinement.c:120,147: trailing soft clip, value = final cursor
            lclip = (first & 15u) == (uint32_t)OP_SOFT;      // refinement.c:210: leading soft clip, keyed by core.pos
        }
        X.rend[r0 + lane] = rendv;
    }
    const unsigned mt = __ballot_sync(FULL, tclip), ml = __ballot_sync(FULL, lclip);
    const unsigned below = (1u << lane) - 1u;
    if (tclip) s.ev[nev + __popc(mt & below)] = make_uint4(rendv, EV_TCLIP, (uint32_t)(r0 + lane), s.endp[lane]);
    if (lclip) s.ev[nev + __popc(mt) + __popc(ml & below)] = make_uint4(s.pos[lane], EV_LCLIP, (uint32_t)(r0 + lane), s.endp[lane]);
    nev += __popc(mt) + __popc(ml);
#if SVTK_IDX_FIXK
    // A chunk whose events all fit its own IDX_FIXK slots (behind the sub-pools) needs no allocation: no atomic round trip at
    // the end of the warp's life, only a fire-and-forget count (second word of the sub-pool's control line) for the totals.
    if (nev > 0 && nev <= IDX_FIXK && head.y == 0u) {
        const uint32_t fb = (uint32_t)IDX_POOLS * X.pool_cap + (uint32_t)c * (uint32_t)IDX_FIXK;
        if (lane == 0) atomicAdd(&X.ctrl[pool * CTRL_STRIDE + 1], (unsigned int)nev);
        __syncwarp();
        const uint32_t packed = index_store_partitioned(s, X, fb, nev);
        head = make_uint4(fb, packed, 0xffffffffu, head.w);
    } else
#endif
    if (nev > 0) index_flush_warp(s, X, pool, nev, head);
    if (lane == 0) X.chunks[c] = head;
}


// ---------------------------------------------------------------------------------------------------------
// Two lower bounds at once over pos[t0,t1) -- first index with pos >= keyA / keyB -- by the whole warp, 32 probes
// per round and both probe loads in flight together (hts.c:3913: yielded reads have pos < end; look-back bound).
// Keys are 64-bit (uint32 window arithmetic widened): out-of-range keys are resolved without probing.
struct WarpSearch {
    int a, b, res;           // open range [a,b); res >= 0 once found
    uint32_t key;
    int idx, step;
    bool fin, ok;
    __device__ __forceinline__ void init(int t0, int t1, long long k) {
        a = t0; b = t1; res = -1; key = (uint32_t)k;
        if (k <= 0) res = t0; else if (k > 0xffffffffll) res = t1;
    }
    __device__ __forceinline__ void probe(int lane) {
        ok = false; fin = false; idx = a; step = 1;
        if (res >= 0) return;
        if (b - a > 32) { step = (b - a + 31) >> 5; idx = min(a + (lane + 1) * step - 1, b - 1); ok = true; }
        else { fin = true; idx = a + lane; ok = idx < b; }
    }
    __device__ __forceinline__ void update(uint32_t v) {
        if (res >= 0) return;
        const unsigned m = __ballot_sync(FULL, ok && v >= key);
        if (fin) res = m ? a + (__ffs(m) - 1) : b;
        else if (!m) res = b;
        else { const int L = __ffs(m) - 1; const int nb = __shfl_sync(FULL, idx, L); a += L * step; b = nb; }
    }
};

__device__ __forceinline__ void warp_lower_bound_pos2(const uint32_t *__restrict__ pos, long long t0, long long t1, long long keyA,
                                                      long long keyB, long long &resA, long long &resB) {
    const int lane = lane_id();
    WarpSearch A, B;
    A.init((int)t0, (int)t1, keyA); B.init((int)t0, (int)t1, keyB);
    while (A.res < 0 || B.res < 0) {
        A.probe(lane); B.probe(lane);
        const uint32_t vA = A.ok ? __ldg(pos + A.idx) : 0u;
        const uint32_t vB = B.ok ? __ldg(pos + B.idx) : 0u;
        A.update(vA); B.update(vB);
    }
    resA = A.res; resB = B.res;
}

// The same two lower bounds through the position tiles (k_pos_tiles): the tile of each key gives the first read at or right of
// the tile's boundary, a warp-wide probe of the next 32 positions finishes the job (a 4 096-base tile holds ~8 reads at 30x).
// Two dependent memory round trips instead of the 32-ary search's four.
__device__ __forceinline__ void warp_lower_bound_tiles2(const DevReads &R, int tid, long long t0, long long t1, long long keyA, long long keyB,
                                                        long long &resA, long long &resB) {
    const int lane = lane_id();
    const long long tb = R.tile_off[tid], nt = R.tile_off[tid + 1] - tb;
    long long a = -1, b = -1;                                   // probe starts (-1: resolved without probing)
    resA = keyA <= 0 ? t0 : t1; resB = keyB <= 0 ? t0 : t1;
    if (keyA > 0 && keyA <= 0xffffffffll && (keyA >> POS_TILE_SHIFT) < nt) a = tb + (keyA >> POS_TILE_SHIFT);
    if (keyB > 0 && keyB <= 0xffffffffll && (keyB >> POS_TILE_SHIFT) < nt) b = tb + (keyB >> POS_TILE_SHIFT);
    const uint32_t sa = a >= 0 ? __ldg(R.tiles + a) : 0u, sb = b >= 0 ? __ldg(R.tiles + b) : 0u;
    long long ia = sa, ib = sb;
    bool needA = a >= 0, needB = b >= 0;
    while (needA || needB) {
        const long long pa = ia + lane, pb = ib + lane;
        const uint32_t va = (needA && pa < t1) ? __ldg(R.pos + pa) : 0xffffffffu;
        const uint32_t vb = (needB && pb < t1) ? __ldg(R.pos + pb) : 0xffffffffu;
        if (needA) {
            const unsigned m = __ballot_sync(FULL, pa >= t1 || (long long)va >= keyA);
            if (m) { resA = min(ia + (__ffs(m) - 1), t1); needA = false; } else ia += 32;
        }
        if (needB) {
            const unsigned m = __ballot_sync(FULL, pb >= t1 || (long long)vb >= keyB);
            if (m) { resB = min(ib + (__ffs(m) - 1), t1); needB = false; } else ib += 32;
        }
    }
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

// What one signature event contributes to a window.  Returns 1 = emit `val`, 2 = leading-clip read whose
// walk stops inside the window (needs warp_stop_cursor), 0 = nothing.
__device__ __forceinline__ int event_for_window(const uint4 ev, int kind, uint32_t w_start, uint32_t w_end, uint32_t beg, long long lo,
                                                long long hi, const uint32_t *__restrict__ rend, int &val) {
    const uint32_t type = ev.y & 15u, len = ev.y >> 4;
    const long long ri = (long long)ev.z;
    if (!(ri >= lo && ri < hi && ev.w > beg)) return 0;      // the iterator yields this read iff endpos > beg (hts.c:3914)
    if (kind == KIND_DEL_START) {
        if (type == EV_DEL) { val = (int)ev.x; return ev.x <= w_end; }
        if (type == EV_TCLIP) { val = (int)ev.x; return w_start <= ev.x && ev.x <= w_end; }
        return 0;
    }
    if (kind == KIND_DEL_END) {
        if (type == EV_DEL) { val = (int)(ev.x + len + 1u); return ev.x <= w_end; }
        if (type == EV_LCLIP && w_start <= ev.x && ev.x <= w_end) {
            const uint32_t rendv = __ldg(rend + ri);
            if (rendv <= w_end) { val = (int)(rendv + 1u); return 1; }   // walk never stopped: final cursor + 1
            val = (int)ev.z;
            return 2;
        }
        return 0;
    }
    if (type == EV_INS) { val = (int)ev.x; return ev.x <= w_end; }       // KIND_INS
    return 0;
}

// ascending bitonic sort of 32*K ints held K per lane, element index = k*32 + lane; no shared memory, no barrier
template <int K>
__device__ __forceinline__ void warp_bitonic_sort(int (&v)[K]) {
    const int lane = lane_id();
#pragma unroll
    for (int size = 2; size <= 32 * K; size <<= 1) {
#pragma unroll
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            if (stride >= 32) {
#pragma unroll
                for (int k = 0; k < K; ++k) {
                    const int kk = k ^ (stride >> 5);
                    if (kk > k) {
                        const bool asc = ((k * 32) & size) == 0;
                        const int lo = min(v[k], v[kk]), hi = max(v[k], v[kk]);
                        v[k] = asc ? lo : hi; v[kk] = asc ? hi : lo;
                    }
                }
            } else {
#pragma unroll
                for (int k = 0; k < K; ++k) {
                    const int o = __shfl_xor_sync(FULL, v[k], stride);
                    const bool asc = (((k * 32) | lane) & size) == 0;
                    const bool lower = (lane & stride) == 0;
                    v[k] = (asc == lower) ? min(v[k], o) : max(v[k], o);
                }
            }
        }
    }
}

template <int K>
__device__ __forceinline__ void warp_sort_segment(int *a, int n) {
    const int lane = lane_id();
    int v[K];
#pragma unroll
    for (int k = 0; k < K; ++k) { const int i = k * 32 + lane; v[k] = i < n ? a[i] : 0x7fffffff; }
    warp_bitonic_sort<K>(v);
#pragma unroll
    for (int k = 0; k < K; ++k) { const int i = k * 32 + lane; if (i < n) a[i] = v[k]; }
    __syncwarp();
}

#ifndef SVTK_RW_WARPS
#define SVTK_RW_WARPS 4
#endif
constexpr int RW_WARPS = SVTK_RW_WARPS;            // windows per CTA of k_refine_warp
constexpr int RW_CAP = 256;            // signatures a warp sorts in registers (8 per lane)
constexpr int RW_DEFER = 8;

struct RefineWarpShared {
    int sig[RW_CAP];
    int defer[RW_DEFER];
};

struct IndexCounters {                // instrumented instantiation only
    unsigned long long events_scanned, segments, pairs_range, signatures, overflow;
};

// ---------------------------------------------------------------------------------------------------------
// One warp per window.  Windows with more than RW_CAP signatures go to the overflow list (window, exact count).
template <bool STATS>
#ifndef SVTK_RW_MINB
#define SVTK_RW_MINB 12
#endif
__global__ void __launch_bounds__(RW_WARPS * 32, SVTK_RW_MINB) k_refine_warp(DevReads R, DevWindows W, DevIndex X, Tunables t, int32_t *__restrict__ result,
                                                               OverflowList ov, int2 *__restrict__ range_out, IndexCounters *ic) {
    __shared__ RefineWarpShared sh[RW_WARPS];
    const int lane = lane_id(), warp = threadIdx.x >> 5;
    const int w = blockIdx.x * RW_WARPS + warp;
    if (w >= W.n_win) return;
    RefineWarpShared &s = sh[warp];
    const int kind = W.kind[w];
    const uint32_t w_start = W.start[w], w_end = W.end[w];
    const int tid = W.chrom[w] - 1;                              // refinement.c:114
    const uint32_t beg = w_start - 1u;
    const long long begq = (long long)beg, endq = (long long)(uint32_t)(w_end - 1u);   // uint32 arithmetic, then widened
    long long lo = 0, hi = 0;
    if (tid >= 0 && tid < R.n_tid && endq >= begq && kind != KIND_POINT) {
        // yielded reads have pos < endq (hts.c:3913) and pos + span > beg; span <= max_span bounds the look-back
        if (R.tiles) warp_lower_bound_tiles2(R, tid, R.tid_off[tid], R.tid_off[tid + 1], endq, begq - (long long)R.max_span + 1, hi, lo);
        else warp_lower_bound_pos2(R.pos, R.tid_off[tid], R.tid_off[tid + 1], endq, begq - (long long)R.max_span + 1, hi, lo);
        if (lo > hi) lo = hi;
    }
    if (lane == 0) range_out[w] = make_int2((int)lo, (int)hi);
    if (hi <= lo) {                                              // no iterator / no reads / refine_point
        if (lane == 0) result[w] = -1;
        return;
    }
    const unsigned below = (1u << lane) - 1u;
    const bool ins_class = kind == KIND_INS;
    // consensus_pos only ever looks at signatures within `range` of pos (plus `interval` for the cluster behind an anchor), at
    // the array's two ends and at its size (refinement.c:43,56-98).  With the values further away replaced by their nearest
    // representative on either side -- the largest one below pos - reach, the smallest one above pos + reach -- every loop of
    // consensus_pos visits the same elements and stops at the same place, so only the near ones are kept and sorted (a window
    // drags in every D/I of every read that overlaps it, SURVEY Q7: ~65 signatures at config 2, ~20 of them near pos).
    // Off when pos +- reach could wrap in the reference's int arithmetic: then everything is kept, as before.
    const int pos = (int)W.pos[w];
    const long long reach = (long long)t.range + (long long)t.interval + 1ll;
    const bool filt = t.range >= 0 && t.interval >= 0 && reach < (1ll << 30) && pos >= 0 && (long long)pos <= 0x7fffffffll - reach - 1ll;
    const long long keep_lo = filt ? (long long)pos - reach : -(1ll << 40), keep_hi = filt ? (long long)pos + reach : (1ll << 40);
    int far_lo = (int)0x80000000, far_hi = 0x7fffffff;           // per lane: largest value below keep_lo, smallest above keep_hi
    int n = 0, nk = 0, nd = 0;                                   // all signatures, the kept ones, deferred leading-clip reads (uniform)
    unsigned any_lo = 0, any_hi = 0;
    auto take = [&](bool emit, int val) {                        // called by all lanes together
        const bool lo_side = emit && (long long)val < keep_lo, hi_side = emit && (long long)val > keep_hi;
        const unsigned m = __ballot_sync(FULL, emit), mk = __ballot_sync(FULL, emit && !lo_side && !hi_side);
        if (emit && !lo_side && !hi_side) { const int slot = nk + __popc(mk & below); if (slot < RW_CAP) s.sig[slot] = val; }
        if (lo_side) far_lo = max(far_lo, val);
        if (hi_side) far_hi = min(far_hi, val);
        any_lo |= __ballot_sync(FULL, lo_side); any_hi |= __ballot_sync(FULL, hi_side);
        n += __popc(m); nk += __popc(mk);
    };
    unsigned long long n_scan = 0, n_seg = 0;
    const long long c0 = lo / IDX_READS, c1 = (hi - 1) / IDX_READS;
    for (long long cb = c0; cb <= c1; cb += 32) {
        // 32 chunk heads at a time, then their events as ONE flat list: lane e takes event e of the concatenation
        uint4 myh = make_uint4(0u, 0u, 0xffffffffu, 0u);
        if (cb + lane <= c1) myh = __ldg(X.chunks + cb + lane);
        seg_class(myh.x, myh.y, ins_class, myh.x, myh.y);            // from here on: base and count of this window's class
        uint32_t incl = myh.y;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const uint32_t o = __shfl_up_sync(FULL, incl, d); if (lane >= d) incl += o; }
        const uint32_t total = __shfl_sync(FULL, incl, 31);
        if (STATS) { n_scan += total; n_seg += (unsigned long long)min(32ll, c1 - cb + 1); }
        for (uint32_t e0 = 0; e0 < total; e0 += 32) {
            const uint32_t e = e0 + lane;
            // owner chunk: first lane whose inclusive count exceeds e (5-step search over the lanes' prefix)
            int j = 0;
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) { const uint32_t v = __shfl_sync(FULL, incl, j + d - 1); if (v <= e) j += d; }
            const uint32_t jbase = __shfl_sync(FULL, myh.x, j), jincl = __shfl_sync(FULL, incl, j), jcnt = __shfl_sync(FULL, myh.y, j);
            int code = 0, val = 0;
            if (e < total)
                code = event_for_window(__ldg(X.events + (size_t)jbase + (e - (jincl - jcnt))), kind, w_start, w_end, beg, lo, hi, X.rend, val);
            take(code == 1, val);
            const unsigned md = __ballot_sync(FULL, code == 2);
            if (md) {
                if (code == 2) { const int slot = nd + __popc(md & below); if (slot < RW_DEFER) s.defer[slot] = val; }
                nd += __popc(md);
            }
        }
        // chunks that flushed more than once chain older segments behind the head (rare): walk those one by one
        unsigned chained = __ballot_sync(FULL, myh.z != 0xffffffffu);
        while (chained) {
            const int j = __ffs(chained) - 1;
            chained &= chained - 1;
            uint32_t nxt = __shfl_sync(FULL, myh.z, j);
            while (nxt != 0xffffffffu) {
                uint4 seg = __ldg(X.segs + nxt);
                seg_class(seg.x, seg.y, ins_class, seg.x, seg.y);
                if (STATS) { n_scan += seg.y; n_seg += 1; }
                for (uint32_t e0 = 0; e0 < seg.y; e0 += 32) {
                    int code = 0, val = 0;
                    if (e0 + lane < seg.y) code = event_for_window(__ldg(X.events + (size_t)seg.x + e0 + lane), kind, w_start, w_end, beg, lo, hi, X.rend, val);
                    take(code == 1, val);
                    const unsigned md = __ballot_sync(FULL, code == 2);
                    if (md) {
                        if (code == 2) { const int slot = nd + __popc(md & below); if (slot < RW_DEFER) s.defer[slot] = val; }
                        nd += __popc(md);
                    }
                }
                nxt = seg.z;
            }
        }
    }
    __syncwarp();
    // leading-clip reads whose walk stops inside the window: value = cursor after the crossing op, + 1 (refinement.c:210-220)
    if (nd > 0) {
        if (nd <= RW_DEFER) {
            for (int k = 0; k < nd; ++k) {
                const long long ri = s.defer[k];
                const uint32_t sr = warp_stop_cursor(R.cigar, __ldg(R.cigar_off + ri), __ldg(R.n_cigar + ri), __ldg(R.pos + ri), w_end);
                take(lane == 0, (int)(sr + 1u));
            }
        } else {
            for (long long ri = lo; ri < hi; ++ri) {             // more than the defer list holds: rescan the read range (rare)
                const uint32_t nc = __ldg(R.n_cigar + ri), rp = __ldg(R.pos + ri);
                if (nc == 0 || !(__ldg(R.end + ri) > beg) || !(w_start <= rp && rp <= w_end)) continue;
                const unsigned long long off = __ldg(R.cigar_off + ri);
                if ((__ldg(R.cigar + off) & 15u) != (uint32_t)OP_SOFT || __ldg(X.rend + ri) <= w_end) continue;
                const uint32_t sr = warp_stop_cursor(R.cigar, off, nc, rp, w_end);
                take(lane == 0, (int)(sr + 1u));
            }
        }
        __syncwarp();
    }
    if (STATS && lane == 0) {
        atomicAdd(&ic->events_scanned, n_scan); atomicAdd(&ic->segments, n_seg);
        atomicAdd(&ic->pairs_range, (unsigned long long)(hi - lo)); atomicAdd(&ic->signatures, (unsigned long long)n);
    }
    if (n < t.min_count || n <= 0) {                             // refinement.c:43: the size test counts every signature
        if (lane == 0) result[w] = -1;
        return;
    }
    // the far values' representatives join the kept ones: [largest below] kept... [smallest above]
    const int nf = nk + (any_lo ? 1 : 0) + (any_hi ? 1 : 0);
    if (nf > RW_CAP) {
        if (lane == 0) {
            const int slot = atomicAdd(ov.count, 1);
            if (slot < ov.capacity) ov.items[slot] = make_int2(w, n);
            result[w] = -1;
            if (STATS) atomicAdd(&ic->overflow, 1ull);
        }
        return;
    }
    if (any_lo) { const int v = __reduce_max_sync(FULL, far_lo); if (lane == 0) s.sig[nk] = v; }
    if (any_hi) { const int v = __reduce_min_sync(FULL, far_hi); if (lane == 0) s.sig[nk + (any_lo ? 1 : 0)] = v; }
    __syncwarp();
    int res;
    if (nf <= 32) {                                              // the common case: sort and both scans in registers
        int v[1] = {lane < nf ? s.sig[lane] : 0x7fffffff};
        warp_bitonic_sort<1>(v);
        res = consensus_pos_lanes(v[0], nf, pos, t);
    } else {
        if (nf <= 64) warp_sort_segment<2>(s.sig, nf);
        else if (nf <= 128) warp_sort_segment<4>(s.sig, nf);
        else warp_sort_segment<8>(s.sig, nf);
        const ScanResult L = consensus_scan_warp(s.sig, nf, pos, t, true);
        if (L.early) res = L.value;
        else {
            const ScanResult Rr = consensus_scan_warp(s.sig, nf, pos, t, false);
            res = Rr.early ? Rr.value : (L.dist < Rr.dist ? L.best : Rr.best);
        }
    }
    if (lane == 0) result[w] = res;
}

// ---------------------------------------------------------------------------------------------------------
// Overflow tier: persistent CTAs walk the list (window, signature count) left by k_refine_warp.  Segment = dynamic shared
// memory when next_pow2(count) fits, else a slice of the global arena (bump-allocated).  Same gather, block-wide.
struct RefineBlockShared {
    int count, n_defer;
    int defer[32];
    unsigned long long base;
    ScanResult res[2];
};

template <int NT>
__global__ void __launch_bounds__(NT) k_refine_block(DevReads R, DevWindows W, DevIndex X, const int2 *__restrict__ range, Tunables t,
                                                     int32_t *__restrict__ result, OverflowList ov, int cap_smem, int *arena,
                                                     unsigned long long arena_ints, unsigned long long *arena_top, int *fail_flag) {
    extern __shared__ int s_dyn[];
    __shared__ RefineBlockShared s;
    constexpr int NWARP = NT / 32;
    const int warp = threadIdx.x >> 5, lane = lane_id();
    const int n_items = min(*ov.count, ov.capacity);
    for (int it = blockIdx.x; it < n_items; it += gridDim.x) {
        const int2 item = ov.items[it];
        const int w = item.x, n_expect = item.y;
        int *seg = s_dyn;
        int cap = cap_smem;
        if (next_pow2(n_expect) > cap_smem) {
            if (threadIdx.x == 0) s.base = atomicAdd(arena_top, (unsigned long long)next_pow2(n_expect));
            __syncthreads();
            if (s.base + (unsigned long long)next_pow2(n_expect) > arena_ints) {     // arena exhausted: host retries with a bigger one
                if (threadIdx.x == 0) atomicExch(fail_flag, 1);
                __syncthreads();
                continue;
            }
            seg = arena + s.base;
            cap = n_expect;
        }
        if (threadIdx.x == 0) { s.count = 0; s.n_defer = 0; }
        __syncthreads();
        const int kind = W.kind[w];
        const uint32_t w_start = W.start[w], w_end = W.end[w], beg = w_start - 1u;
        const long long lo = range[w].x, hi = range[w].y;
        const SigSink sink{seg, &s.count, cap};
        const long long c0 = lo / IDX_READS, c1 = (hi - 1) / IDX_READS;
        for (long long c = c0 + warp; c <= c1; c += NWARP) {
            uint4 sg = __ldg(X.chunks + c);
            while (true) {
                uint32_t base, cnt;
                seg_class(sg.x, sg.y, kind == KIND_INS, base, cnt);
                for (uint32_t e0 = 0; e0 < cnt; e0 += 32) {
                    int code = 0, val = 0;
                    if (e0 + lane < cnt) code = event_for_window(__ldg(X.events + (size_t)base + e0 + lane), kind, w_start, w_end, beg, lo, hi, X.rend, val);
                    if (code == 2) { const int k = atomicAdd(&s.n_defer, 1); if (k < 32) s.defer[k] = val; }
                    const unsigned m = __ballot_sync(FULL, code == 1);
                    if (m) {
                        int b = 0;
                        if (lane == 0) b = atomicAdd(&s.count, __popc(m));
                        b = __shfl_sync(FULL, b, 0);
                        if (code == 1) { const int slot = b + __popc(m & ((1u << lane) - 1u)); if (slot < cap) seg[slot] = val; }
                    }
                }
                if (sg.z == 0xffffffffu) break;
                sg = __ldg(X.segs + sg.z);
            }
        }
        __syncthreads();
        const int nd = s.n_defer;
        if (nd > 0) {
            if (nd <= 32) {
                for (int k = warp; k < nd; k += NWARP) {
                    const long long ri = s.defer[k];
                    const uint32_t sr = warp_stop_cursor(R.cigar, __ldg(R.cigar_off + ri), __ldg(R.n_cigar + ri), __ldg(R.pos + ri), w_end);
                    if (lane == 0) sink.push((int)(sr + 1u));
                }
            } else {
                for (long long ri = lo + warp; ri < hi; ri += NWARP) {
                    const uint32_t nc = __ldg(R.n_cigar + ri), rp = __ldg(R.pos + ri);
                    if (nc == 0 || !(__ldg(R.end + ri) > beg) || !(w_start <= rp && rp <= w_end)) continue;
                    const unsigned long long off = __ldg(R.cigar_off + ri);
                    if ((__ldg(R.cigar + off) & 15u) != (uint32_t)OP_SOFT || __ldg(X.rend + ri) <= w_end) continue;
                    const uint32_t sr = warp_stop_cursor(R.cigar, off, nc, rp, w_end);
                    if (lane == 0) sink.push((int)(sr + 1u));
                }
            }
            __syncthreads();
        }
        __threadfence_block();
        const int n = s.count;                  // == n_expect
        block_bitonic_sort(seg, n);
        const int res = block_consensus(seg, n, (int)W.pos[w], t, s.res);
        if (threadIdx.x == 0) result[w] = res;
        __syncthreads();
    }
}

// svtk_reads contract check: CIGARs packed back to back in read order (cigar_off[r] + n_cigar[r] == cigar_off[r+1]) -> bit 0;
// reads sorted by pos inside every reference -> bit 1; end >= pos -> bit 2
__global__ void k_check_layout(DevReads R, unsigned long long n_words, unsigned int *flag) {
    for (long long r = blockIdx.x * (long long)blockDim.x + threadIdx.x; r < R.n_reads; r += (long long)gridDim.x * blockDim.x) {
        const unsigned long long next = (r + 1 < R.n_reads) ? R.cigar_off[r + 1] : n_words;
        unsigned bad = 0;
        if (R.cigar_off[r] + R.n_cigar[r] != next || (r == 0 && R.cigar_off[0] != 0ull)) bad |= 1u;
        if (R.end[r] < R.pos[r]) bad |= 4u;
        if (r > 0 && R.pos[r] < R.pos[r - 1]) {               // allowed only across a reference boundary
            int lo = 0, hi = R.n_tid;                          // first t with tid_off[t] > r - 1
            while (lo < hi) { const int m = (lo + hi) >> 1; if (R.tid_off[m] > r - 1) hi = m; else lo = m + 1; }
            if (lo >= R.n_tid + 1 || R.tid_off[lo] != r) bad |= 2u;
        }
        if (bad) atomicOr(flag, bad);
    }
}

// position of the last read of every reference (0xffffffff: the reference has no reads)
__global__ void k_last_pos(DevReads R, uint32_t *__restrict__ out) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= R.n_tid) return;
    const long long a = R.tid_off[t], z = R.tid_off[t + 1];
    out[t] = z > a ? R.pos[z - 1] : 0xffffffffu;
}

// position tiles: tiles[tile_off[t] + k] = first read of reference t with pos >= k << POS_TILE_SHIFT (reference end if none);
// one thread per read writes the tiles whose boundary falls in (pos[r-1], pos[r]], the reference's last read also the tail
__global__ void k_pos_tiles(DevReads R, uint32_t *__restrict__ tiles, const long long *__restrict__ tile_off) {
    const long long r = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (r >= R.n_reads) return;
    int lo = 0, hi = R.n_tid;                                  // reference of read r: last t with tid_off[t] <= r
    while (lo < hi) { const int m = (lo + hi) >> 1; if (R.tid_off[m + 1] > r) hi = m; else lo = m + 1; }
    const int t = lo;
    const long long t0 = R.tid_off[t], t1 = R.tid_off[t + 1], tb = tile_off[t], nt = tile_off[t + 1] - tb;
    // tiles k with (k << SHIFT) in (pos[r-1], pos[r]]: k from floor(pos[r-1] / T) + 1 (a boundary AT pos[r-1] belongs to read r-1 or earlier)
    const long long first = (r == t0) ? 0 : (long long)(R.pos[r - 1] >> POS_TILE_SHIFT) + 1;
    const long long mine = (long long)(R.pos[r] >> POS_TILE_SHIFT);
    for (long long k = first; k <= mine && k < nt; ++k) tiles[tb + k] = (uint32_t)r;
    if (r + 1 == t1) for (long long k = mine + 1; k < nt; ++k) tiles[tb + k] = (uint32_t)t1;
}

// =========================================================================================================
constexpr int RB_NT = 1024;

int index_reads_per_chunk() { return IDX_READS; }
int index_pools() { return IDX_POOLS; }
int index_zeroes_next_ctrl() { return 1; }
int index_fixed_slots() { return SVTK_IDX_FIXK; }
int index_ctrl_words() { return CTRL_ERR + CTRL_STRIDE; }
int index_ctrl_stride() { return CTRL_STRIDE; }
int refine_block_segment_ints() { return 32768; }      // power of two: the bitonic sort pads a segment to next_pow2(n)

cudaError_t launch_check_layout(const LaunchReads &r, unsigned long long n_words, unsigned int *flag, cudaStream_t s) {
    if (r.n_reads <= 0) return cudaSuccess;
    const long long want = (r.n_reads + 255) / 256;
    const int blocks = (int)(want < 148 * 8 ? want : 148 * 8);
    k_check_layout<<<blocks, 256, 0, s>>>(to_dev(r), n_words, flag);
    return cudaGetLastError();
}

int pos_tile_shift() { return POS_TILE_SHIFT; }

cudaError_t launch_last_pos(const LaunchReads &r, uint32_t *out, cudaStream_t s) {
    if (r.n_tid <= 0) return cudaSuccess;
    k_last_pos<<<(r.n_tid + 127) / 128, 128, 0, s>>>(to_dev(r), out);
    return cudaGetLastError();
}

cudaError_t launch_pos_tiles(const LaunchReads &r, uint32_t *tiles, const long long *tile_off, cudaStream_t s) {
    if (r.n_reads <= 0) return cudaSuccess;
    k_pos_tiles<<<(unsigned)((r.n_reads + 255) / 256), 256, 0, s>>>(to_dev(r), tiles, tile_off);
    return cudaGetLastError();
}

cudaError_t launch_index_reads(const LaunchReads &r, const LaunchIndex &x, cudaStream_t s) {
    if (r.n_reads <= 0) return cudaSuccess;
    k_index_reads_defer<<<(x.n_chunks + IDX_WARPS - 1) / IDX_WARPS, IDX_WARPS * 32, 0, s>>>(to_dev(r), to_dev(x));
    return cudaGetLastError();
}

cudaError_t launch_refine_indexed(const LaunchReads &r, const LaunchWindows &w, const LaunchIndex &x, int2 *range, int range_, int interval,
                                  int min_count, int32_t *result, int *ov_count, int2 *ov_items, int ov_capacity,
                                  unsigned long long *counters, cudaStream_t s) {
    if (w.n_win <= 0) return cudaSuccess;
    const Tunables t{range_, interval, min_count};
    const OverflowList ov{ov_count, ov_items, ov_capacity};
    const int grid = (w.n_win + RW_WARPS - 1) / RW_WARPS;
    if (counters)
        k_refine_warp<true><<<grid, RW_WARPS * 32, 0, s>>>(to_dev(r), to_dev(w), to_dev(x), t, result, ov, range,
                                                            reinterpret_cast<IndexCounters *>(counters));
    else
        k_refine_warp<false><<<grid, RW_WARPS * 32, 0, s>>>(to_dev(r), to_dev(w), to_dev(x), t, result, ov, range, nullptr);
    return cudaGetLastError();
}

cudaError_t launch_refine_block(const LaunchReads &r, const LaunchWindows &w, const LaunchIndex &x, const int2 *range, int range_,
                                int interval, int min_count, int32_t *result, int *ov_count, int2 *ov_items, int ov_capacity, int n_items,
                                int *arena, unsigned long long arena_ints, unsigned long long *arena_top, int *fail_flag, cudaStream_t s) {
    static unsigned long long attr_seen = 0;
    const int cap = refine_block_segment_ints();
    if (first_use_on_device(attr_seen)) {
        cudaError_t e = cudaFuncSetAttribute(k_refine_block<RB_NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, cap * 4);
        if (e != cudaSuccess) return e;
    }
    const Tunables t{range_, interval, min_count};
    const OverflowList ov{ov_count, ov_items, ov_capacity};
    const int grid = n_items < 148 ? n_items : 148;
    k_refine_block<RB_NT><<<grid, RB_NT, cap * 4, s>>>(to_dev(r), to_dev(w), to_dev(x), range, t, result, ov, cap, arena, arena_ints,
                                                       arena_top, fail_flag);
    return cudaGetLastError();
}

}  // namespace svtk
