// tools/ab/index_variants.cuh -- NOT COMPILED.  Two superseded forms of the signature-index pass, kept for the record of the A/B
// measurements in DESIGN.md 7 (B200, config 2, us per launch): k_index_reads (events resolved inside every step) 41.7;
// k_index_reads_lane (one lane per read, direct or TMA-staged) 51.5 / 68; the product kernel k_index_reads_defer
// (svtrek_b200/csrc/svtk_index.cu) 35.1.  They were compiled from inside svtk_index.cu (same helpers, same DevIndex layout) under
// SVTK_IDX_DEFER=0 / SVTK_IDX_LANE=1 until round 2; to rebuild one, paste it back above k_index_reads_defer.

// ---------------------------------------------------------------------------------------------------------
constexpr int IDX_EVCAP = IDX_TILE;           // events a warp stages in shared memory between flushes (>= one step's worst case)
struct IndexWarpShared {
    uint32_t P[IDX_TILE];                // exclusive cursor prefix (chunk relative) of every word of the step
    uint4 ev[IDX_EVCAP];
    uint32_t rel[IDX_READS + 4];         // word offset of each read, relative to the aligned chunk start; 0xffffffff beyond nr
    uint32_t pos[IDX_READS], endp[IDX_READS];
    uint32_t base[IDX_READS + 1];        // P at each read's first word
    uint2 optab[16];                     // per CIGAR op: {advance multiplier, smallest word that is a signature}
};

#ifndef SVTK_IDX_MINB
#define SVTK_IDX_MINB 8
#endif
__global__ void __launch_bounds__(IDX_WARPS * 32, SVTK_IDX_MINB) k_index_reads(DevReads R, DevIndex X) {
    __shared__ __align__(16) IndexWarpShared sh[IDX_WARPS];
    const int lane = lane_id(), warp = threadIdx.x >> 5;
    const long long c = (long long)blockIdx.x * IDX_WARPS + warp;
    if (c >= X.n_chunks) return;
    IndexWarpShared &s = sh[warp];
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
        // x = 2^28 for ops that advance the reference cursor (all but I and S, refinement.c:137): hi32(word * 2^28) == len.
        // y = first word value that is a signature: I with len >= 50 (refinement.c:299), D with len > 50 (refinement.c:124).
        const uint32_t adv = (lane == OP_INS || lane == OP_SOFT) ? 0u : (1u << 28);
        const uint32_t thr = lane == OP_INS ? ((uint32_t)SV_MIN_LENGTH << 4 | OP_INS)
                           : lane == OP_DEL ? ((uint32_t)(SV_MIN_LENGTH + 1) << 4 | OP_DEL) : 0xffffffffu;
        s.optab[lane] = make_uint2(adv, thr);
    }
    __syncwarp();
    const uint32_t head_pad = (uint32_t)(w0 - w0a);
    const uint32_t span = s.rel[nr];
    const uint32_t *__restrict__ cg = R.cigar + w0a;
    uint4 head = make_uint4(0u, 0u, 0xffffffffu, s.pos[0]);

    uint32_t run = 0;       // cursor advance of all words before the current step (uniform)
    int nev = 0;            // staged events (uniform)
#if SVTK_IDX_PREFETCH
    // software pipeline: the next step's 32 bytes per lane are in flight while the current step is decoded
    uint4 nx[IDX_WPT / 4];
#pragma unroll
    for (int q = 0; q < IDX_WPT / 4; ++q) {
        nx[q] = make_uint4(0u, 0u, 0u, 0u);
        if ((uint32_t)lane * IDX_WPT + 4u * q < span) nx[q] = __ldg(reinterpret_cast<const uint4 *>(cg + lane * IDX_WPT + 4 * q));
    }
#endif
    for (uint32_t tb = 0; tb < span; tb += IDX_TILE) {
        const uint32_t i0 = tb + (uint32_t)lane * IDX_WPT;
        uint32_t w[IDX_WPT];
#pragma unroll
        for (int q = 0; q < IDX_WPT / 4; ++q) {
#if SVTK_IDX_PREFETCH
            w[4 * q] = nx[q].x; w[4 * q + 1] = nx[q].y; w[4 * q + 2] = nx[q].z; w[4 * q + 3] = nx[q].w;
            nx[q] = make_uint4(0u, 0u, 0u, 0u);
            if (i0 + IDX_TILE + 4u * q < span) nx[q] = __ldg(reinterpret_cast<const uint4 *>(cg + i0 + IDX_TILE + 4u * q));
#else
            uint4 v = make_uint4(0u, 0u, 0u, 0u);
            if (i0 + 4u * q < span) v = __ldg(reinterpret_cast<const uint4 *>(cg + i0 + 4u * q));
            w[4 * q] = v.x; w[4 * q + 1] = v.y; w[4 * q + 2] = v.z; w[4 * q + 3] = v.w;
#endif
        }
        if (tb == 0 || tb + IDX_TILE > span) {               // edge steps: blank the neighbouring chunks' words ("0M", inert)
#pragma unroll
            for (int k = 0; k < IDX_WPT; ++k) { const uint32_t gi = i0 + k; if (gi < head_pad || gi >= span) w[k] = 0u; }
        }
        uint32_t p[IDX_WPT];
        uint32_t sum = 0;
        unsigned candm = 0;
#pragma unroll
        for (int k = 0; k < IDX_WPT; ++k) {
#if SVTK_IDX_TABLE
            const uint2 ot = s.optab[w[k] & 15u];            // shared-memory lookup: keeps the ALU pipe for the compares
            p[k] = sum;
            sum = mad_hi_u32(w[k], ot.x, sum);               // += len for ops that advance
            if (w[k] >= ot.y) candm |= 1u << k;
#else
            // 16-entry bit tables indexed by the op nibble (rotate by the low 5 bits of the word; the tables have period 16)
            const uint32_t keep = __funnelshift_r(0xFFEDFFEDu, 0xFFEDFFEDu, w[k]) & 1u;   // every op except I(1), S(4) advances
            const uint32_t indel = __funnelshift_r(0x00060006u, 0x00060006u, w[k]) & 1u;  // op is I or D
            p[k] = sum;
            sum += keep * (w[k] >> 4);
            if (indel && w[k] >= 0x321u && w[k] != 0x322u) candm |= 1u << k;
#endif
        }
        const uint32_t incl = warp_inclusive_scan(sum);
        const uint32_t mybase = run + (incl - sum);
#pragma unroll
        for (int q = 0; q < IDX_WPT / 4; ++q)
            *reinterpret_cast<uint4 *>(&s.P[lane * IDX_WPT + 4 * q]) =
                make_uint4(mybase + p[4 * q], mybase + p[4 * q + 1], mybase + p[4 * q + 2], mybase + p[4 * q + 3]);
        const int myc = __popc(candm);
        const int tot = __reduce_add_sync(FULL, myc);        // candidates of this step (uniform)
        __syncwarp();
        if (lane < nr) {
            const uint32_t rel = s.rel[lane];
            if (rel >= tb && rel < tb + IDX_TILE && rel < span) s.base[lane] = s.P[rel - tb];
        }
        if (tot) {
            if (nev + tot > IDX_EVCAP) { index_flush_warp(s, X, pool, nev, head); nev = 0; }
            __syncwarp();
            int slot = myc;                                  // exclusive prefix of the per-lane candidate counts
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) { const int o = __shfl_up_sync(FULL, slot, d); if (lane >= d) slot += o; }
            slot += nev - myc;
            unsigned cm = candm;
            while (cm) {
                const int k = __ffs(cm) - 1;
                cm &= cm - 1;
                const uint32_t i = (uint32_t)(lane * IDX_WPT + k), gi = tb + i;
                const uint4 ra = *reinterpret_cast<const uint4 *>(&s.rel[0]), rb = *reinterpret_cast<const uint4 *>(&s.rel[4]);
                // the read owning word gi: last r with rel[r] <= gi (rel[0] <= gi always; entries beyond nr are 0xffffffff)
                const int r = (int)(ra.y <= gi) + (int)(ra.z <= gi) + (int)(ra.w <= gi) + (int)(rb.x <= gi) + (int)(rb.y <= gi) +
                              (int)(rb.z <= gi) + (int)(rb.w <= gi);
                const uint32_t word = __ldg(cg + gi);
                const uint32_t x = s.pos[r] + (s.P[i] - s.base[r]);
                const uint32_t type = ((word & 15u) == (uint32_t)OP_DEL) ? EV_DEL : EV_INS;
                s.ev[slot++] = make_uint4(x, (word & ~15u) | type, (uint32_t)(r0 + r), s.endp[r]);
            }
            nev += tot;
        }
        run += __shfl_sync(FULL, incl, 31);
        __syncwarp();                                        // P is rewritten by the next step
    }
    // reads that start at the very end of the chunk's span (empty CIGARs) and the end-of-span sentinel
    if (lane <= nr && s.rel[lane] >= span) s.base[lane] = run;
    __syncwarp();
    if (nev + 2 * IDX_READS > IDX_EVCAP) { index_flush_warp(s, X, pool, nev, head); nev = 0; }
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
    if (nev > 0) index_flush_warp(s, X, pool, nev, head);
    if (lane == 0) X.chunks[c] = head;
}


#if SVTK_IDX_LANE
// ---------------------------------------------------------------------------------------------------------
// k_index_reads_lane: the same index, built with ONE LANE PER READ.  A lane walks its own read's CIGAR front to back
// with 128-bit loads, so the reference cursor is a plain running sum in a register: no warp scan, no per-word prefix
// array, no read-boundary bookkeeping -- about 6 instructions per CIGAR word instead of ~46 (ncu r01f: 1.44 warp
// instructions per word in k_index_reads).  A warp owns 32 consecutive reads = 4 chunks of IDX_READS; every lane stages
// its events in a private shared-memory column (LANE_EV slots) and the warp flushes them per chunk with one atomic per
// chunk when a column fills up and at the end.  Loads are lane-strided (each lane streams its own 16-byte pieces), so the
// L1 processes one line per lane and load: the price for dropping the scans, paid in LSU cycles instead of issue slots.
constexpr int LANE_WARPS = 2;
constexpr int LANE_EV = 8;                  // staged events per lane between flushes
constexpr int LANE_STAGE_WORDS = 5120;      // CIGAR words of the warp's 32 reads staged in shared memory (20 KB; 32 x 150-word HiFi reads = 4800)

struct __align__(128) LaneWarpShared {
    uint32_t stage[LANE_STAGE_WORDS];       // the warp's contiguous CIGAR span, brought in by ONE bulk (TMA) copy
    uint4 ev[LANE_EV * 32];                 // slot s of lane l at [s * 32 + l]: conflict-free 128-bit stores
    uint2 optab[16];
    unsigned long long mbar;                // completion barrier of the bulk copy
};

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

// flush the staged events of the warp's four chunks; head = newest segment of this lane's chunk (uniform inside the 8-lane group)
__device__ __forceinline__ void lane_flush(LaneWarpShared &s, const DevIndex &X, int pool, int &cnt, uint4 &head) {
    const int lane = lane_id(), sl = lane & 7;
    int incl = cnt;
#pragma unroll
    for (int d = 1; d < 8; d <<= 1) { const int t = __shfl_up_sync(FULL, incl, d, 8); if (sl >= d) incl += t; }
    const int subtot = __shfl_sync(FULL, incl, 7, 8);
    const int excl = incl - cnt;
    uint32_t fb = 0xffffffffu;
    uint4 nh = head;
    if (sl == 0 && subtot > 0) {
        const unsigned int b = atomicAdd(&X.ctrl[pool * CTRL_STRIDE], (unsigned int)subtot);
        if ((unsigned long long)b + (unsigned long long)subtot > (unsigned long long)X.pool_cap) atomicOr(&X.ctrl[CTRL_ERR], 1u);
        else {
            uint32_t next = 0xffffffffu;
            bool ok = true;
            if (head.y != 0u) {                              // push the previous segment down the chain
                const unsigned int si = atomicAdd(&X.ctrl[CTRL_SEG], 1u);
                if (si >= X.seg_cap) { atomicOr(&X.ctrl[CTRL_ERR], 2u); ok = false; }
                else { X.segs[si] = head; next = si; }
            }
            if (ok) { fb = (uint32_t)pool * X.pool_cap + b; nh = make_uint4(fb, (uint32_t)subtot, next, head.w); }
        }
    }
    fb = __shfl_sync(FULL, fb, 0, 8);
    nh.x = __shfl_sync(FULL, nh.x, 0, 8); nh.y = __shfl_sync(FULL, nh.y, 0, 8); nh.z = __shfl_sync(FULL, nh.z, 0, 8);
    head = nh;
    if (fb != 0xffffffffu)
        for (int i = 0; i < cnt; ++i) X.events[(size_t)fb + excl + i] = s.ev[i * 32 + lane];
    cnt = 0;
    __syncwarp();
}

__global__ void __launch_bounds__(LANE_WARPS * 32) k_index_reads_lane(DevReads R, DevIndex X) {
    extern __shared__ __align__(128) unsigned char lane_smem[];
    LaneWarpShared *sh = reinterpret_cast<LaneWarpShared *>(lane_smem);
    const int lane = lane_id(), warp = threadIdx.x >> 5;
    const long long r0 = ((long long)blockIdx.x * LANE_WARPS + warp) * 32;
    if (r0 >= R.n_reads) return;
    LaneWarpShared &s = sh[warp];
    const int pool = (int)(blockIdx.x % IDX_POOLS);
    if (lane < 16) {
        const uint32_t adv = (lane == OP_INS || lane == OP_SOFT) ? 0u : (1u << 28);
        const uint32_t thr = lane == OP_INS ? ((uint32_t)SV_MIN_LENGTH << 4 | OP_INS)
                           : lane == OP_DEL ? ((uint32_t)(SV_MIN_LENGTH + 1) << 4 | OP_DEL) : 0xffffffffu;
        s.optab[lane] = make_uint2(adv, thr);
    }
    __syncwarp();
    const long long r = r0 + lane;
    const bool have = r < R.n_reads;
    unsigned long long off = 0;
    uint32_t n = 0, rpos = 0, endp = 0;
    if (have) { off = __ldg(R.cigar_off + r); n = __ldg(R.n_cigar + r); rpos = __ldg(R.pos + r); endp = __ldg(R.end + r); }
    const unsigned long long al = off & ~3ull;
    const uint32_t headw = (uint32_t)(off - al);
    const uint32_t nv = n ? (headw + n + 3u) >> 2 : 0u;      // 16-byte pieces this lane reads
    // The 32 reads' CIGARs are one contiguous run of words (svtk_reads contract).  One lane issues a single bulk copy
    // (cp.async.bulk: the TMA unit streams it, fully coalesced) into the warp's stage; the lanes then read their own
    // reads from shared memory, where lane-strided 128-bit loads cost bank conflicts instead of one L1 line each.
    // A run longer than the stage (long-CIGAR reads) is walked straight from global memory.
    const unsigned long long span0 = __shfl_sync(FULL, al, 0);
    const int last_lane = (int)min((long long)31, R.n_reads - 1 - r0);
    const unsigned long long span1 = (__shfl_sync(FULL, off + n, last_lane) + 3ull) & ~3ull;
    const bool staged = span1 - span0 <= (unsigned long long)LANE_STAGE_WORDS;
    const uint4 *cg4 = reinterpret_cast<const uint4 *>(R.cigar + al);
    if (staged && span1 > span0) {
        const uint32_t bar = smem_u32(&s.mbar), bytes = (uint32_t)(span1 - span0) * 4u;
        if (lane == 0) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         ::"r"(smem_u32(s.stage)), "l"(R.cigar + span0), "r"(bytes), "r"(bar) : "memory");
        }
        __syncwarp();
        uint32_t done = 0;
        while (!done)
            asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0, 1, 0, p; }" : "=r"(done) : "r"(bar) : "memory");
        cg4 = reinterpret_cast<const uint4 *>(s.stage + (al - span0));
    }
    const uint32_t nvmax = __reduce_max_sync(FULL, nv);
    uint4 head = make_uint4(0u, 0u, 0xffffffffu, __shfl_sync(FULL, rpos, 0, 8));
    uint32_t cur = rpos;
    int cnt = 0;
    uint4 nxt = make_uint4(0u, 0u, 0u, 0u);
    if (nv > 0) nxt = cg4[0];
    for (uint32_t v = 0; v < nvmax; ++v) {
        if (v < nv) {
            uint32_t w[4] = {nxt.x, nxt.y, nxt.z, nxt.w};
            if (v + 1 < nv) nxt = cg4[v + 1];
            if (v == 0 || v + 1 == nv) {                     // neighbours' words in the first / last piece: "0M", inert
#pragma unroll
                for (int k = 0; k < 4; ++k) { const uint32_t gi = 4u * v + k; if (gi < headw || gi >= headw + n) w[k] = 0u; }
            }
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const uint2 ot = s.optab[w[k] & 15u];
                if (w[k] >= ot.y) {                          // D > 50 / I >= 50 (refinement.c:124,299): cursor BEFORE the op
                    const uint32_t type = ((w[k] & 15u) == (uint32_t)OP_DEL) ? EV_DEL : EV_INS;
                    s.ev[cnt * 32 + lane] = make_uint4(cur, (w[k] & ~15u) | type, (uint32_t)r, endp);
                    ++cnt;
                }
                cur = mad_hi_u32(w[k], ot.x, cur);           // += len for every op except I and S (refinement.c:137)
            }
        }
        if (__any_sync(FULL, cnt > LANE_EV - 4)) lane_flush(s, X, pool, cnt, head);   // room for one more piece
    }
    // soft-clip events and the final cursor
    if (__any_sync(FULL, cnt > LANE_EV - 2)) lane_flush(s, X, pool, cnt, head);
    if (have) {
        if (n > 0) {
            const uint32_t first = __ldg(R.cigar + off), last = __ldg(R.cigar + off + n - 1);
            if ((last & 15u) == (uint32_t)OP_SOFT) { s.ev[cnt * 32 + lane] = make_uint4(cur, EV_TCLIP, (uint32_t)r, endp); ++cnt; }     // refinement.c:120,147
            if ((first & 15u) == (uint32_t)OP_SOFT) { s.ev[cnt * 32 + lane] = make_uint4(rpos, EV_LCLIP, (uint32_t)r, endp); ++cnt; }   // refinement.c:210
        }
        X.rend[r] = cur;
    }
    lane_flush(s, X, pool, cnt, head);
    const long long c = r0 / IDX_READS + (lane >> 3);
    if ((lane & 7) == 0 && c < X.n_chunks) X.chunks[c] = head;
}

#endif  // SVTK_IDX_LANE
