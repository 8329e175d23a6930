// svtk_ins.cu -- svtk_ins_sequences: the inserted bases behind every refined insertion breakpoint.
//
// The reference stops at the breakpoint: refine_ins (refinement.c:278-325) pushes the reference position of every `I >= 50`
// op and consensus_pos picks a cluster; the bases are never looked at, and abPOA -- built into the binary -- is never called
// (SURVEY F5).  The north star names "abPOA for insertion consensus", so this entry produces what that call needs: for each
// INS window whose refinement succeeded, the inserted sequence of every `I >= 50` op that refine_ins would have pushed
// (same reads: sam_itr_queryi's selection, hts.c:3913-3916; same walk: cursor advanced by every op except I and S, stop once
// it passes W.end, refinement.c:299-318) and that lies within `tol` bases of the refined position, as one read set per
// window in svtk_poa_consensus' layout (abPOA codes A0 C1 G2 T3 other 4).  Sequences come from the BAM image
// svtk_load_bam left in HBM: 4-bit SEQ (bam_get_seq / bam_seqi, htslib/sam.h) at the query offset of the op.
//
//   k_ins_hits    one CTA per INS window, one warp per candidate read: warp scans over the CIGAR keep the reference cursor
//                 and the query offset; qualifying ops are counted (pass 1) or appended to the hit list (pass 2)
//   k_ins_bases   one warp per hit: nibbles -> codes, written at the offset the host laid out after sorting the hits by
//                 (window, read, query offset) -- the order a reader of the file would meet them
#include "svtk_ctx.h"

#include <cuda_runtime.h>

#include <algorithm>
#include <cstring>
#include <vector>

namespace svtk {

struct InsHit {
    int32_t win;
    int32_t read;
    uint32_t ref_pos, qoff, len;
    uint32_t pad;
};

constexpr unsigned FULLM = 0xffffffffu;

__device__ __forceinline__ long long lower_bound_u32(const uint32_t *__restrict__ a, long long lo, long long hi, long long key) {
    while (lo < hi) {                                       // first index with a[i] >= key (key may be negative / beyond u32)
        const long long mid = (lo + hi) >> 1;
        if ((long long)a[mid] < key) lo = mid + 1; else hi = mid;
    }
    return lo;
}

__device__ __forceinline__ uint32_t warp_incl_scan(uint32_t v, int lane) {
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const uint32_t o = __shfl_up_sync(FULLM, v, d); if (lane >= d) v += o; }
    return v;
}

__global__ void __launch_bounds__(128) k_ins_hits(LaunchReads R, LaunchWindows W, const int32_t *__restrict__ refined, int tol, int min_len,
                                                  const uint32_t *__restrict__ l_seq, InsHit *__restrict__ hits, unsigned long long cap,
                                                  unsigned long long *__restrict__ counter) {
    const int w = blockIdx.x;
    if (W.kind[w] != 2 || refined[w] < 0) return;            // SVTK_INS windows whose refinement gave a position
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t w_start = W.start[w], w_end = W.end[w];
    const int tid = W.chrom[w] - 1;                          // refinement.c:114
    const uint32_t beg = w_start - 1u;
    const long long begq = (long long)beg, endq = (long long)(uint32_t)(w_end - 1u);   // uint32 arithmetic, then widened
    if (tid < 0 || tid >= R.n_tid || endq < begq) return;
    const long long t0 = R.tid_off[tid], t1 = R.tid_off[tid + 1];
    const long long hi = lower_bound_u32(R.pos, t0, t1, endq);                          // yielded reads have pos < endq (hts.c:3913)
    long long lo = lower_bound_u32(R.pos, t0, t1, begq - (long long)R.max_span + 1);    // and pos + span > beg
    if (lo > hi) lo = hi;
    const long long P = (long long)refined[w];
    for (long long r = lo + warp; r < hi; r += 4) {
        if (!(R.end[r] > beg)) continue;                     // hts.c:3914
        const uint32_t n = R.n_cigar[r];
        const unsigned long long off = R.cigar_off[r];
        uint32_t cur = R.pos[r], q = 0;
        const uint32_t ls = l_seq[r];
        for (uint32_t b = 0; b < n; b += 32) {
            const uint32_t cw = (b + lane < n) ? R.cigar[off + b + lane] : 0u;
            const uint32_t op = cw & 15u, len = cw >> 4;
            const uint32_t adv = ((0x12u >> op) & 1u) ? 0u : len;              // every op but I(1) and S(4) advances (refinement.c:312)
            const uint32_t qadv = ((0x193u >> op) & 1u) ? len : 0u;            // M I S = X consume the query (bam_cigar_type bit 0)
            const uint32_t ia = warp_incl_scan(adv, lane), iq = warp_incl_scan(qadv, lane);
            const uint32_t before = cur + (ia - adv), qbefore = q + (iq - qadv);
            const long long d = (long long)before - P;
            const bool hit = (b + lane < n) && op == 1u && len >= (uint32_t)min_len && before <= w_end && d >= -(long long)tol && d <= (long long)tol &&
                             (unsigned long long)qbefore + len <= (unsigned long long)ls;   // SEQ `*` (l_seq 0): nothing to extract
            if (hit) {
                const unsigned long long k = atomicAdd(counter, 1ull);
                if (k < cap) { InsHit h; h.win = w; h.read = (int32_t)r; h.ref_pos = before; h.qoff = qbefore; h.len = len; h.pad = 0; hits[k] = h; }
            }
            cur += __shfl_sync(FULLM, ia, 31); q += __shfl_sync(FULLM, iq, 31);
            if (cur > w_end) break;                          // later ops start past W.end: the reference has stopped (refinement.c:314)
        }
    }
}

__device__ __forceinline__ uint32_t ld8u(const uint32_t *__restrict__ w, unsigned long long off) {
    return (w[off >> 2] >> ((uint32_t)(off & 3) * 8u)) & 255u;
}

__global__ void __launch_bounds__(256) k_ins_bases(const uint32_t *__restrict__ infl, const unsigned long long *__restrict__ seq_src,
                                                   const InsHit *__restrict__ hits, const unsigned long long *__restrict__ out_off, long long n_hits,
                                                   uint8_t *__restrict__ bases) {
    const long long h = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (h >= n_hits) return;
    const InsHit x = hits[h];
    const unsigned long long src = seq_src[x.read], dst = out_off[h];
    for (uint32_t j = lane; j < x.len; j += 32) {
        const uint32_t qi = x.qoff + j;
        const uint32_t byte = ld8u(infl, src + (qi >> 1));
        const uint32_t nib = (qi & 1u) ? (byte & 15u) : (byte >> 4);           // bam_seqi: high nibble first
        bases[dst + j] = nib == 1u ? 0 : nib == 2u ? 1 : nib == 4u ? 2 : nib == 8u ? 3 : 4;   // =ACMGRSVTWYHKDBN -> abPOA's A0 C1 G2 T3 other 4
    }
}

}  // namespace svtk

using namespace svtk;

#define CU(call)                                                                                              \
    do {                                                                                                      \
        cudaError_t e_ = (call);                                                                              \
        if (e_ != cudaSuccess) return fail(SVTK_ERR_CUDA, "%s: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

extern "C" int svtk_ins_sequences(svtk_ctx *ctx, const svtk_windows *w, const int32_t *refined, int32_t tol, int32_t min_len, svtk_ins_sets *out) {
    Ctx *c = reinterpret_cast<Ctx *>(ctx);
    if (!c || !w || !out || (w->n_win > 0 && !refined)) return fail(SVTK_ERR_ARG, "null argument");
    memset(out, 0, sizeof(*out));
    if (!c->have_reads || !c->reads_from_bam)
        return fail(SVTK_ERR_STATE, "svtk_ins_sequences needs the BAM image: load the reads with svtk_load_bam");
    if (tol < 0 || min_len < 1) return fail(SVTK_ERR_ARG, "bad tolerance / minimum length");
    if (int rc = set_device(c)) return rc;
    if (c->pending) if (int rc = svtk_sync(ctx)) return rc;
    c->ins_set_off.assign(1, 0); c->ins_seq_off.assign(1, 0); c->ins_bases.clear(); c->ins_window.clear(); c->ins_read.clear(); c->ins_ref_pos.clear();
    auto publish = [&]() {
        out->n_sets = (int32_t)c->ins_window.size(); out->n_seqs = (int64_t)c->ins_read.size();
        out->set_off = c->ins_set_off.data(); out->seq_off = c->ins_seq_off.data(); out->bases = c->ins_bases.data();
        out->window = c->ins_window.data(); out->read = c->ins_read.data(); out->ref_pos = c->ins_ref_pos.data();
    };
    const size_t n = (size_t)w->n_win;
    if (n == 0) { publish(); return 0; }
    // windows + refined positions -> device (same blob layout as svtk_refine_windows)
    const size_t n4 = (n * 4 + 15) & ~(size_t)15, blob = 5 * n4 + ((n + 15) & ~(size_t)15);
    if (int rc = c->h_win.reserve(blob)) return rc;
    if (int rc = c->win_blob.reserve(blob)) return rc;
    char *hb = (char *)c->h_win.p;
    memcpy(hb, w->chrom, 4 * n); memcpy(hb + n4, w->start, 4 * n); memcpy(hb + 2 * n4, w->end, 4 * n);
    memcpy(hb + 3 * n4, w->pos, 4 * n); memcpy(hb + 4 * n4, refined, 4 * n); memcpy(hb + 5 * n4, w->kind, n);
    CU(cudaMemcpyAsync(c->win_blob.p, hb, blob, cudaMemcpyHostToDevice, c->stream));
    char *db = (char *)c->win_blob.p;
    const LaunchWindows lw{w->n_win, (const int32_t *)db, (const uint32_t *)(db + n4), (const uint32_t *)(db + 2 * n4),
                           (const uint32_t *)(db + 3 * n4), (const uint8_t *)(db + 5 * n4)};
    const int32_t *d_ref = (const int32_t *)(db + 4 * n4);
    const size_t nr = (size_t)c->reads.n_reads;
    const unsigned long long *seq_src = (const unsigned long long *)c->bam_rec.p + 2 * nr;
    const uint32_t *l_seq = (const uint32_t *)(seq_src + nr);
    if (int rc = c->bam_ctl.reserve(256)) return rc;
    unsigned long long *counter = (unsigned long long *)((char *)c->bam_ctl.p + 128);
    unsigned long long n_hits = 0;
    // pass 1 counts, pass 2 fills a list of exactly that size
    CU(cudaMemsetAsync(counter, 0, 8, c->stream));
    k_ins_hits<<<(unsigned)n, 128, 0, c->stream>>>(c->reads, lw, d_ref, tol, min_len, l_seq, nullptr, 0ull, counter);
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(&n_hits, counter, 8, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    c->launches += 1;
    if (n_hits == 0) { publish(); return 0; }
    if (n_hits >= (1ull << 31)) return fail(SVTK_ERR_ARG, "more than 2^31 inserted sequences in one batch: split the windows");
    if (int rc = c->tmp_a.reserve(sizeof(InsHit) * (size_t)n_hits)) return rc;
    CU(cudaMemsetAsync(counter, 0, 8, c->stream));
    k_ins_hits<<<(unsigned)n, 128, 0, c->stream>>>(c->reads, lw, d_ref, tol, min_len, l_seq, (InsHit *)c->tmp_a.p, n_hits, counter);
    CU(cudaGetLastError());
    std::vector<InsHit> hits((size_t)n_hits);
    CU(cudaMemcpyAsync(hits.data(), c->tmp_a.p, sizeof(InsHit) * (size_t)n_hits, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    c->launches += 1;
    std::sort(hits.begin(), hits.end(), [](const InsHit &a, const InsHit &b) {
        return a.win != b.win ? a.win < b.win : a.read != b.read ? a.read < b.read : a.qoff < b.qoff;
    });
    std::vector<unsigned long long> off((size_t)n_hits);
    unsigned long long total = 0;
    for (size_t i = 0; i < hits.size(); ++i) {
        if (i == 0 || hits[i].win != hits[i - 1].win) { if (i) c->ins_set_off.push_back((int64_t)i); c->ins_window.push_back(hits[i].win); }
        off[i] = total; total += hits[i].len;
        c->ins_seq_off.push_back((int64_t)total); c->ins_read.push_back(hits[i].read); c->ins_ref_pos.push_back(hits[i].ref_pos);
    }
    c->ins_set_off.push_back((int64_t)n_hits);
    // sorted hits + output offsets back to the device, bases out
    if (int rc = c->tmp_b.reserve(8 * (size_t)n_hits)) return rc;
    if (int rc = c->tmp_c.reserve((size_t)total + 16)) return rc;
    if (int rc = upload(c, c->tmp_a.p, hits.data(), sizeof(InsHit) * (size_t)n_hits)) return rc;
    if (int rc = upload(c, c->tmp_b.p, off.data(), 8 * (size_t)n_hits)) return rc;
    const long long threads = (long long)n_hits * 32;
    k_ins_bases<<<(unsigned)((threads + 255) / 256), 256, 0, c->stream>>>((const uint32_t *)c->bam_infl.p, seq_src, (const InsHit *)c->tmp_a.p,
                                                                          (const unsigned long long *)c->tmp_b.p, (long long)n_hits, (uint8_t *)c->tmp_c.p);
    CU(cudaGetLastError());
    c->ins_bases.resize((size_t)total);
    CU(cudaMemcpyAsync(c->ins_bases.data(), c->tmp_c.p, (size_t)total, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    c->launches += 1;
    c->d2h_bytes += (size_t)total + sizeof(InsHit) * (size_t)n_hits;
    publish();
    return 0;
}
