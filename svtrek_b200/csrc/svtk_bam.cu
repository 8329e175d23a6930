// svtk_bam.cu -- svtk_load_bam: a coordinate-sorted, BAI-indexed BAM file -> the resident read set, decoded on the GPU.
//
// Replaces, for the audt path, what the reference does through HTSlib on every refine_* call: hts_open / sam_index_load
// (audit.c:270-272), sam_itr_queryi + sam_itr_next (refinement.c:109-117 and the three siblings), i.e. bgzf_read_block ->
// inflate (bgzf.c:988, 746), bam_read1 (sam.c:754-795, incl. the CG-tag long-CIGAR fix-up sam.c:645) and bam_endpos
// (sam.c:638-643).  The reference inflates the blocks under every window again and again (>= 99 % of its wall time,
// SURVEY.md 6); here the file crosses PCIe once, compressed, and
//
//   k_bgzf_inflate   every BGZF block at once: one LANE per block (32 independent DEFLATE streams per warp, tables in
//                    shared memory, svtk_bgzf_core.cuh), persistent CTAs pulling block numbers from one counter
//   k_bam_walk       record boundaries: BAM records form a chain (each starts where the previous one's block_size ends), so
//                    the chain is cut at the record-aligned virtual offsets the .bai already holds (linear-index entries and
//                    bin chunk starts); one thread walks each piece -- first to count reads / CIGAR words, then to fill
//   k_bam_extract    one warp per record: core fields, CIGAR words copied into the packed array, bam_endpos
//   k_bam_tid_off    read ranges per reference id + the coordinate-order check
//
// and the result is adopted exactly like svtk_attach_reads_device() would (same arrays, same invariants).
// Every record with refID >= 0 becomes a read, whatever its flags (the reference filters nothing).
#include "svtk_ctx.h"
#include "svtk_bgzf_core.cuh"

#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <memory>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

namespace svtk {

struct BgzfBlock {
    unsigned long long in_beg;     // byte offset of the DEFLATE payload inside the file image
    unsigned long long out_off;    // byte offset of the block's first inflated byte
    uint32_t in_len, isize;
};

enum { BAM_CTL_NEXT = 0, BAM_CTL_ZERR = 1, BAM_CTL_ZCODE = 2, BAM_CTL_WERR = 4, BAM_CTL_WSEG = 5, BAM_CTL_ORDER = 6, BAM_CTL_CRC = 7 };
enum { BW_OK = 0, BW_ERR_SHORT = 1, BW_ERR_SIZE = 2, BW_ERR_FIELDS = 3, BW_ERR_AUX = 4, BW_ERR_END = 5 };

constexpr unsigned FULLMASK = 0xffffffffu;

// ---------------------------------------------------------------------------------------------------------
// `lanes` = lanes per warp that decode (1 .. 32): the decoder is bound by the instruction latency of one warp, and a warp pays
// for the union of its lanes' paths (literal / match / long code / block switch), so a file with few blocks runs faster
// spread over many thin warps; a large file needs all 32 lanes to fit the machine.
#ifndef SVTK_BZ_MINB
#define SVTK_BZ_MINB 16           // resident single-warp CTAs per SM the register allocation must allow
#endif
__global__ void __launch_bounds__(32, SVTK_BZ_MINB) k_bgzf_inflate(const uint32_t *__restrict__ words, const BgzfBlock *__restrict__ blk, unsigned first_block,
                                                     unsigned end_block, uint8_t *__restrict__ out, unsigned long long *__restrict__ tok,
                                                     unsigned long long tok_sub, uint32_t *__restrict__ n_tok, unsigned int *__restrict__ ctl,
                                                     unsigned int *__restrict__ next, int lanes) {
    extern __shared__ uint32_t bz_sm[];
    // length / distance bases (svtk_bgzf_core.cuh): one copy per CTA, behind the lanes' Huffman tables
    uint32_t *lut = bz_sm + lanes * BZ_TABLE_WORDS;
    if (threadIdx.x < 29u) lut[threadIdx.x] = bz_len_entry((int)threadIdx.x);
    if (threadIdx.x < 30u) lut[32 + threadIdx.x] = bz_dist_entry((int)threadIdx.x);
    __syncwarp();
    const BzLut L{lut};
    BzTables T;
    T.lit = reinterpret_cast<uint16_t *>(bz_sm + (threadIdx.x < (unsigned)lanes ? threadIdx.x : 0u) * BZ_TABLE_WORDS);
    T.dist = T.lit + (1 << BZ_LBITS);
    BzLane z;
    z.state = BZ_S_IDLE; z.err = BZ_OK; z.n_tok = 0; z.tok = nullptr;
    unsigned cur = 0xffffffffu;
    bool drained = threadIdx.x >= (unsigned)lanes;
    for (;;) {
        if (z.state == BZ_S_IDLE && !drained) {
            if (cur != 0xffffffffu) {
                n_tok[cur] = z.err ? 0u : z.n_tok;              // a failed block has nothing to resolve
                if (z.err) { if (atomicCAS(&ctl[BAM_CTL_ZERR], 0u, cur + 1u) == 0u) ctl[BAM_CTL_ZCODE] = (unsigned)z.err; }
            }
            cur = first_block + atomicAdd(next, 1u);
            if (cur < end_block) {
                const BgzfBlock b = blk[cur];
                bz_begin(z, words, b.in_beg, b.in_beg + b.in_len, out + b.out_off, b.isize, tok + (bz_tok_offset(b.out_off, cur) - tok_sub));
            } else { drained = true; z.err = BZ_OK; cur = 0xffffffffu; }
        }
        if (__all_sync(FULLMASK, drained)) break;
        // every lane passes here once per round: block switches (header + table build) on the lanes that need one, then a
        // burst of symbol rounds that all lanes take together
        if (z.state == BZ_S_HEADER) bz_header(z, T);
#pragma unroll 1
        for (int it = 0; it < BZ_BURST; ++it)
            if (z.state == BZ_S_SYMBOLS) bz_step(z, T, L);
        bz_check_input(z);
    }
}

// Second phase: the matches of every block, in stream order, one WARP per block, 32 tokens at a time.  A match may be copied
// once every byte of its source is final: always true for the lowest pending match of the batch (everything before its
// destination is final), and for any other whose source ends below that match's write position (the "frontier").  Each round
// every ready lane copies one chunk of BZ_CHUNK bytes; the lowest pending lane always progresses, so the loop ends.
__global__ void __launch_bounds__(256) k_bgzf_resolve(const BgzfBlock *__restrict__ blk, unsigned first_block, unsigned end_block, uint8_t *__restrict__ out,
                                                      const unsigned long long *__restrict__ tok, unsigned long long tok_sub, const uint32_t *__restrict__ n_tok) {
    const unsigned lane = threadIdx.x & 31u, warps = gridDim.x * (blockDim.x >> 5);
    for (unsigned b = first_block + blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); b < end_block; b += warps) {
        const uint32_t n = n_tok[b];
        if (n == 0u) continue;
        const unsigned long long ob = blk[b].out_off;
        uint8_t *base = out + ob;
        const unsigned long long *tq = tok + (bz_tok_offset(ob, b) - tok_sub);
        for (uint32_t k0 = 0; k0 < n; k0 += 32u) {
            const bool have = k0 + lane < n;
            const unsigned long long t = have ? tq[k0 + lane] : 0ull;
            uint32_t wr = bz_tok_pos(t), left = have ? bz_tok_len(t) : 0u, dist = bz_tok_dist(t);    // wr: next byte this lane writes
            const uint32_t src_end = wr - dist + left;                                               // one past the last source byte
            unsigned pm = __ballot_sync(FULLMASK, left != 0u);
            while (pm) {
                const int low = __ffs(pm) - 1;
                const uint32_t frontier = __shfl_sync(FULLMASK, wr, low);
                if (left != 0u && ((int)lane == low || src_end <= frontier)) {
                    const uint32_t c = left < BZ_CHUNK ? left : BZ_CHUNK;
                    bz_copy_chunk(base + wr, &dist, c);
                    wr += c; left -= c;
                }
                __syncwarp();                                      // this round's stores are visible to the next round's loads
                pm = __ballot_sync(FULLMASK, left != 0u);
            }
        }
    }
}

// CRC-32 of every block's inflated bytes against its gzip footer (HTSlib: bgzf.c:776-780 "CRC32 checksum mismatch").  One warp
// per block, 32 slices, merged with the x^(8n) identity (svtk_bgzf_core.cuh); word loads once a lane's slice is aligned.
__global__ void __launch_bounds__(256) k_bgzf_crc(const uint32_t *__restrict__ comp, const BgzfBlock *__restrict__ blk, unsigned n_blocks,
                                                  const uint8_t *__restrict__ infl, unsigned int *__restrict__ ctl) {
    __shared__ uint32_t tab[256];
    tab[threadIdx.x] = bz_crc_table_entry(threadIdx.x);
    __syncthreads();
    const unsigned lane = threadIdx.x & 31u, warps = gridDim.x * (blockDim.x >> 5);
    for (unsigned b = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); b < n_blocks; b += warps) {
        const BgzfBlock k = blk[b];
        const uint32_t n = k.isize, per = ((n + 31u) / 32u + 3u) & ~3u;
        const unsigned long long s0 = (unsigned long long)lane * per;
        const uint32_t len = s0 >= n ? 0u : min(per, n - (uint32_t)s0);
        const uint8_t *p = infl + k.out_off + s0;
        uint32_t c = 0xffffffffu, i = 0;
        for (; i < len && ((uintptr_t)(p + i) & 3u); ++i) c = tab[(c ^ p[i]) & 255u] ^ (c >> 8);
        for (; i + 4u <= len; i += 4u) {
            uint32_t w = *reinterpret_cast<const uint32_t *>(p + i);
#pragma unroll
            for (int q = 0; q < 4; ++q) { c = tab[(c ^ w) & 255u] ^ (c >> 8); w >>= 8; }
        }
        for (; i < len; ++i) c = tab[(c ^ p[i]) & 255u] ^ (c >> 8);
        uint32_t share = (len > 0u || lane == 0u) ? bz_crc_share(c ^ 0xffffffffu, lane, per, n) : 0u;
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) share ^= __shfl_xor_sync(FULLMASK, share, d);
        if (lane == 0u) {
            const unsigned long long f = k.in_beg + k.in_len;                  // gzip footer: CRC32, ISIZE
            const uint32_t want = bz_byte(comp, f) | bz_byte(comp, f + 1) << 8 | bz_byte(comp, f + 2) << 16 | bz_byte(comp, f + 3) << 24;
            if (share != want) atomicCAS(&ctl[BAM_CTL_CRC], 0u, b + 1u);
        }
    }
}

// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t ld32(const uint32_t *__restrict__ w, unsigned long long off) {
    const unsigned long long i = off >> 2;
    return __funnelshift_r(w[i], w[i + 1], (uint32_t)(off & 3) * 8u);
}
__device__ __forceinline__ uint32_t ld8(const uint32_t *__restrict__ w, unsigned long long off) {
    return (w[off >> 2] >> ((uint32_t)(off & 3) * 8u)) & 255u;
}

// bam_aux_get(b, "CG") restricted to what bam_tag2cigar accepts (sam.c:645-672): returns the byte offset of the first real
// CIGAR word and its count, or 0 words when the record keeps its in-place CIGAR; -1 on malformed aux data
__device__ int find_cg(const uint32_t *__restrict__ w, unsigned long long p, unsigned long long end, uint32_t n_cigar, unsigned long long *src) {
    while (p + 3 <= end) {
        const uint32_t t0 = ld8(w, p), t1 = ld8(w, p + 1), ty = ld8(w, p + 2);
        p += 3;
        unsigned long long sz;
        if (ty == 'A' || ty == 'c' || ty == 'C') sz = 1;
        else if (ty == 's' || ty == 'S') sz = 2;
        else if (ty == 'i' || ty == 'I' || ty == 'f') sz = 4;
        else if (ty == 'd') sz = 8;
        else if (ty == 'Z' || ty == 'H') {
            unsigned long long q = p;
            while (q < end && ld8(w, q) != 0u) ++q;
            if (q >= end) return -1;
            sz = q + 1 - p;
        } else if (ty == 'B') {
            if (p + 5 > end) return -1;
            const uint32_t sub = ld8(w, p), cnt = ld32(w, p + 1);
            const uint32_t es = (sub == 'c' || sub == 'C') ? 1u : (sub == 's' || sub == 'S') ? 2u : (sub == 'i' || sub == 'I' || sub == 'f') ? 4u : 0u;
            if (es == 0u) return -1;
            if (t0 == 'C' && t1 == 'G') {
                if ((sub == 'I' || sub == 'i') && cnt >= n_cigar && cnt < (1u << 29) && p + 5 + 4ull * cnt <= end) { *src = p + 5; return (int)cnt; }
                return 0;
            }
            sz = 5 + (unsigned long long)es * cnt;
        } else return -1;
        if (t0 == 'C' && t1 == 'G') return 0;                // a CG tag that is not B,I: CIGAR untouched
        if (p + sz > end) return -1;
        p += sz;
    }
    return p == end ? 0 : -1;
}

// One thread per piece [splits[2s], splits[2s+1]) of the record chain (pieces of different loaded ranges do not touch).  FILL == false: seg[s] = {reads, CIGAR words}.
// FILL == true: seg[s] holds the exclusive prefix {first read index, first CIGAR word}; per read the record offset, the
// CIGAR source offset, n_cigar and cigar_off are written.
template <bool FILL>
__global__ void __launch_bounds__(128) k_bam_walk(const uint32_t *__restrict__ w, const unsigned long long *__restrict__ splits, int n_seg,
                                                  ulonglong2 *__restrict__ seg, unsigned long long *__restrict__ rec_off,
                                                  unsigned long long *__restrict__ cig_src, uint32_t *__restrict__ n_cigar_out,
                                                  unsigned long long *__restrict__ cigar_off, unsigned long long *__restrict__ seq_src,
                                                  uint32_t *__restrict__ l_seq_out, unsigned int *__restrict__ ctl) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_seg) return;
    unsigned long long off = splits[2 * s];
    const unsigned long long end = splits[2 * s + 1];
    unsigned long long nr = 0, nw = 0;
    if (FILL) { const ulonglong2 b = seg[s]; nr = b.x; nw = b.y; }
    int err = BW_OK;
    while (off < end) {
        if (off + 36 > end) { err = BW_ERR_SHORT; break; }
        const uint32_t bs = ld32(w, off);
        const int32_t refid = (int32_t)ld32(w, off + 4);
        if (bs < 32u || off + 4ull + bs > end) { err = BW_ERR_SIZE; break; }
        if (refid >= 0) {
            const int32_t pos = (int32_t)ld32(w, off + 8);
            const uint32_t lrn = ld32(w, off + 12) & 255u, w4 = ld32(w, off + 16), l_seq = ld32(w, off + 20);
            uint32_t nc = w4 & 0xffffu;
            unsigned long long src = off + 36ull + lrn;
            const unsigned long long seq_at = src + 4ull * nc;             // SEQ follows the CIGAR the record stores in place
            const unsigned long long aux = src + 4ull * nc + ((unsigned long long)l_seq + 1) / 2 + l_seq, rend = off + 4ull + bs;
            if (aux > rend) { err = BW_ERR_FIELDS; break; }
            if (nc > 0u && pos >= 0) {
                const uint32_t c0 = ld32(w, src);
                if ((c0 & 15u) == 4u && (c0 >> 4) == l_seq && aux < rend) {          // placeholder <l_seq>S...: look for CG:B,I
                    unsigned long long cg = 0;
                    const int k = find_cg(w, aux, rend, nc, &cg);
                    if (k < 0) { err = BW_ERR_AUX; break; }
                    if (k > 0) { nc = (uint32_t)k; src = cg; }
                }
            }
            if (FILL) { rec_off[nr] = off; cig_src[nr] = src; n_cigar_out[nr] = nc; cigar_off[nr] = nw; seq_src[nr] = seq_at; l_seq_out[nr] = l_seq; }
            ++nr; nw += nc;
        }
        off += 4ull + bs;
    }
    if (!err && off != end) err = BW_ERR_END;
    if (err) { if (atomicCAS(&ctl[BAM_CTL_WERR], 0u, (unsigned)err) == 0u) ctl[BAM_CTL_WSEG] = (unsigned)s; }
    if (!FILL) seg[s] = make_ulonglong2(nr, nw);
}

// one warp per read: pos, tid, CIGAR copy, bam_endpos (sam.c:638-643: FUNMAP or zero reference length -> pos + 1)
__global__ void __launch_bounds__(256) k_bam_extract(const uint32_t *__restrict__ w, const unsigned long long *__restrict__ rec_off,
                                                     const unsigned long long *__restrict__ cig_src, const uint32_t *__restrict__ n_cigar,
                                                     const unsigned long long *__restrict__ cigar_off, long long n_reads, uint32_t *__restrict__ pos_out,
                                                     uint32_t *__restrict__ end_out, int32_t *__restrict__ tid_out, uint32_t *__restrict__ cigar) {
    const long long r = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (r >= n_reads) return;
    const unsigned long long off = rec_off[r], src = cig_src[r], dst = cigar_off[r];
    const uint32_t nc = n_cigar[r];
    uint32_t rlen = 0;
    for (uint32_t i = lane; i < nc; i += 32) {
        const uint32_t cw = ld32(w, src + 4ull * i);
        cigar[dst + i] = cw;
        if ((0x3C1A7u >> ((cw & 15u) << 1)) & 2u) rlen += cw >> 4;      // bam_cigar_type bit 1: consumes reference (sam.h BAM_CIGAR_TYPE)
    }
    rlen = __reduce_add_sync(FULLMASK, rlen);
    if (lane == 0) {
        const uint32_t pos = ld32(w, off + 8), flag = ld32(w, off + 16) >> 16;
        if ((flag & 4u) || rlen == 0u) rlen = 1u;
        pos_out[r] = pos; end_out[r] = pos + rlen; tid_out[r] = (int32_t)ld32(w, off + 4);
    }
}

// tid_off[t] = first read of reference t (reads are in (tid, pos) order); thread n_reads closes the table
__global__ void k_bam_tid_off(const int32_t *__restrict__ tid, const uint32_t *__restrict__ pos, long long n_reads, int n_tid,
                              long long *__restrict__ tid_off, unsigned int *__restrict__ ctl) {
    const long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (r > n_reads) return;
    const int t = r < n_reads ? tid[r] : n_tid, tp = r > 0 ? tid[r - 1] : -1;
    if (r < n_reads && (t >= n_tid || t < tp || (t == tp && (int32_t)pos[r] < (int32_t)pos[r - 1]))) { atomicOr(&ctl[BAM_CTL_ORDER], 1u); return; }
    for (int x = tp + 1; x <= t; ++x) tid_off[x] = r;
}

}  // namespace svtk

// ---------------------------------------------------------------------------------------------------------
using namespace svtk;

#define CU(call)                                                                                              \
    do {                                                                                                      \
        cudaError_t e_ = (call);                                                                              \
        if (e_ != cudaSuccess) return fail(SVTK_ERR_CUDA, "%s: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

namespace {

struct Mapped {                      // read-only mmap of a whole file (the .bai)
    const uint8_t *p = nullptr;
    size_t n = 0;
    int fd = -1;
    int open_file(const char *path) {
        fd = ::open(path, O_RDONLY);
        if (fd < 0) return -1;
        struct stat st;
        if (fstat(fd, &st) != 0) return -1;
        n = (size_t)st.st_size;
        if (n == 0) return 0;
        void *m = mmap(nullptr, n, PROT_READ, MAP_PRIVATE, fd, 0);
        if (m == MAP_FAILED) return -1;
        p = (const uint8_t *)m;
        return 0;
    }
    ~Mapped() {
        if (p) munmap((void *)p, n);
        if (fd >= 0) ::close(fd);
    }
};

inline uint32_t le16(const uint8_t *p) { return (uint32_t)p[0] | (uint32_t)p[1] << 8; }
inline uint32_t le32(const uint8_t *p) { return (uint32_t)p[0] | (uint32_t)p[1] << 8 | (uint32_t)p[2] << 16 | (uint32_t)p[3] << 24; }
inline uint64_t le64(const uint8_t *p) { return (uint64_t)le32(p) | (uint64_t)le32(p + 4) << 32; }

bool pread_all(int fd, void *dst, size_t n, size_t off) {
    char *d = (char *)dst;
    while (n) {
        const ssize_t k = pread(fd, d, n, (off_t)off);
        if (k <= 0) return false;
        d += k; n -= (size_t)k; off += (size_t)k;
    }
    return true;
}

// one BGZF block header (SAMv1 4.1: gzip member with the BC extra subfield carrying BSIZE) at h[0..avail): 0 = ok
int parse_bgzf_header(const uint8_t *h, size_t avail, size_t file_off, uint32_t *bsize, uint32_t *xlen_out) {
    if (avail < 18) return fail(SVTK_ERR_IO, "bam: truncated BGZF header at byte %zu", file_off);
    if (h[0] != 0x1f || h[1] != 0x8b || h[2] != 8 || !(h[3] & 4)) return fail(SVTK_ERR_IO, "bam: not a BGZF block at byte %zu", file_off);
    const uint32_t xlen = le16(h + 10);
    if (12 + (size_t)xlen > avail) return fail(SVTK_ERR_IO, "bam: truncated BGZF extra field at byte %zu", file_off);
    uint32_t bs = 0;
    for (uint32_t x = 0; x + 4 <= xlen;) {
        const uint8_t *e = h + 12 + x;
        const uint32_t slen = le16(e + 2);
        if (e[0] == 'B' && e[1] == 'C' && slen == 2 && x + 6 <= xlen) bs = le16(e + 4) + 1;
        x += 4 + slen;
    }
    if (bs < xlen + 20) return fail(SVTK_ERR_IO, "bam: bad BGZF block size at byte %zu", file_off);
    *bsize = bs; *xlen_out = xlen;
    return 0;
}

// BGZF block table of f[from, n) (used for the header's own blocks): scanning stops at the first block starting at or after `to`
int scan_bgzf(const uint8_t *f, size_t n, size_t from, size_t to, std::vector<BgzfBlock> &blk) {
    size_t off = from;
    unsigned long long out = blk.empty() ? 0ull : blk.back().out_off + blk.back().isize;
    while (off < n && off < to) {
        uint32_t bsize = 0, xlen = 0;
        if (int rc = parse_bgzf_header(f + off, std::min<size_t>(n - off, 65536), off, &bsize, &xlen)) return rc;
        if (off + bsize > n) return fail(SVTK_ERR_IO, "bam: bad BGZF block size at byte %zu", off);
        BgzfBlock b;
        b.in_beg = off + 12 + xlen; b.in_len = bsize - xlen - 20; b.isize = le32(f + off + bsize - 4); b.out_off = out;
        if (b.isize > 65536u) return fail(SVTK_ERR_IO, "bam: BGZF block at byte %zu claims %u inflated bytes", off, b.isize);
        blk.push_back(b);
        out += b.isize;
        off += bsize;
    }
    return 0;
}

// the BAM header (magic, text, reference dictionary) out of the first bytes of the file: inflated on the host with the same
// decoder, block by block, until complete.  *need_more: the buffer ended inside the header (the caller reads more and retries).
int read_header(const uint8_t *f, size_t n, bool whole_file, std::vector<std::string> &names, std::vector<uint32_t> &lens,
                unsigned long long *first_rec, bool *need_more) {
    std::vector<BgzfBlock> blk;                           // the header's own blocks, scanned lazily from the start of the file
    size_t scanned = 0;
    std::vector<uint32_t> words;                          // the decoder reads aligned words: one block's payload at a time
    std::vector<uint8_t> hd;
    size_t nb = 0;
    std::vector<uint16_t> tabs((1 << BZ_LBITS) + (1 << BZ_DBITS));
    BzTables T;
    T.lit = tabs.data(); T.dist = tabs.data() + (1 << BZ_LBITS);
    *need_more = false;
    auto more = [&](size_t need) -> int {
        while (hd.size() < need) {
            if (nb >= blk.size()) {
                // the next block must lie wholly inside the buffer
                if (scanned + 18 > n) { *need_more = !whole_file; return fail(SVTK_ERR_IO, "bam: header runs past the end of the file"); }
                uint32_t bsize = 0, xlen = 0;
                if (int rc = parse_bgzf_header(f + scanned, n - scanned, scanned, &bsize, &xlen)) return rc;
                if (scanned + bsize > n) { *need_more = !whole_file; return fail(SVTK_ERR_IO, "bam: header runs past the end of the file"); }
                if (int rc = scan_bgzf(f, n, scanned, scanned + 1, blk)) return rc;
                scanned += bsize;
            }
            const BgzfBlock &b = blk[nb++];
            words.assign((size_t)(b.in_len + BZ_OVERREAD) / 4 + 4, 0u);
            memcpy(words.data(), f + b.in_beg, b.in_len);
            const size_t at = hd.size();
            hd.resize(at + b.isize);
            uint8_t dummy = 0;
            std::vector<unsigned long long> tok((size_t)b.isize / 3 + 2);
            const int rc = bz_inflate(words.data(), 0, b.in_len, b.isize ? hd.data() + at : &dummy, b.isize, T, tok.data());
            if (rc) return fail(SVTK_ERR_IO, "bam: BGZF block %zu does not inflate (code %d)", nb - 1, rc);
        }
        return 0;
    };
    if (int rc = more(12)) return rc;
    if (memcmp(hd.data(), "BAM\1", 4) != 0) return fail(SVTK_ERR_IO, "bam: bad magic");
    const uint32_t l_text = le32(hd.data() + 4);
    size_t p = 8 + (size_t)l_text;
    if (int rc = more(p + 4)) return rc;
    const int32_t n_ref = (int32_t)le32(hd.data() + p);
    p += 4;
    if (n_ref < 0) return fail(SVTK_ERR_IO, "bam: negative n_ref");
    names.clear(); lens.clear();
    for (int i = 0; i < n_ref; ++i) {
        if (int rc = more(p + 4)) return rc;
        const uint32_t ln = le32(hd.data() + p);
        p += 4;
        if (ln == 0 || ln > (1u << 20)) return fail(SVTK_ERR_IO, "bam: bad reference name length");
        if (int rc = more(p + ln + 4)) return rc;
        names.emplace_back((const char *)hd.data() + p, strnlen((const char *)hd.data() + p, ln));
        p += ln;
        lens.push_back(le32(hd.data() + p));
        p += 4;
    }
    *first_rec = p;
    return 0;
}

// ---- the .bai (SAMv1 5.2), parsed once per load
struct BaiBin { uint32_t bin; uint32_t first, count; };    // chunks [first, first + count) of BaiRefIdx::chunks
struct BaiRefIdx {
    std::vector<BaiBin> bins;                              // sorted by bin number
    std::vector<uint64_t> chunks;                          // pairs (begin, end) virtual offsets
    std::vector<uint64_t> lin;                             // linear index, 16 kb tiles
    uint64_t beg = ~0ull, end = 0;                         // first record / past the last record of the reference
};
struct Bai {
    std::vector<BaiRefIdx> refs;
    std::vector<uint64_t> voffs;                           // every record-aligned offset the index knows (sorted, unique): cut points
};

int read_bai(const char *path, Bai &bai) {
    Mapped m;
    if (m.open_file(path) != 0) return fail(SVTK_ERR_IO, "bam: cannot open index %s (the reference needs it too: sam_index_load, audit.c:272)", path);
    const uint8_t *f = m.p;
    const size_t n = m.n;
    if (n < 8 || memcmp(f, "BAI\1", 4) != 0) return fail(SVTK_ERR_IO, "bam: %s is not a BAI index", path);
    const int32_t n_ref = (int32_t)le32(f + 4);
    if (n_ref < 0) return fail(SVTK_ERR_IO, "bam: bad index");
    bai.refs.assign((size_t)n_ref, BaiRefIdx());
    size_t p = 8;
    for (int r = 0; r < n_ref; ++r) {
        BaiRefIdx &R = bai.refs[(size_t)r];
        if (p + 4 > n) return fail(SVTK_ERR_IO, "bam: truncated index");
        const int32_t n_bin = (int32_t)le32(f + p);
        p += 4;
        for (int b = 0; b < n_bin; ++b) {
            if (p + 8 > n) return fail(SVTK_ERR_IO, "bam: truncated index");
            const uint32_t bin = le32(f + p);
            const int32_t n_chunk = (int32_t)le32(f + p + 4);
            p += 8;
            if (n_chunk < 0 || p + 16ull * (size_t)n_chunk > n) return fail(SVTK_ERR_IO, "bam: truncated index");
            if (bin != 37450u) {                                    // the pseudo-bin holds file-level metadata, not record offsets
                R.bins.push_back(BaiBin{bin, (uint32_t)(R.chunks.size() / 2), (uint32_t)n_chunk});
                for (int k = 0; k < n_chunk; ++k) {
                    const uint64_t vb = le64(f + p + 16 * (size_t)k), ve = le64(f + p + 16 * (size_t)k + 8);
                    R.chunks.push_back(vb); R.chunks.push_back(ve);
                    bai.voffs.push_back(vb);
                    R.beg = std::min(R.beg, vb); R.end = std::max(R.end, ve);
                }
            }
            p += 16 * (size_t)n_chunk;
        }
        std::sort(R.bins.begin(), R.bins.end(), [](const BaiBin &a, const BaiBin &b) { return a.bin < b.bin; });
        if (p + 4 > n) return fail(SVTK_ERR_IO, "bam: truncated index");
        const int32_t n_intv = (int32_t)le32(f + p);
        p += 4;
        if (n_intv < 0 || p + 8ull * (size_t)n_intv > n) return fail(SVTK_ERR_IO, "bam: truncated index");
        R.lin.resize((size_t)n_intv);
        for (int k = 0; k < n_intv; ++k) { const uint64_t v = le64(f + p + 8 * (size_t)k); R.lin[(size_t)k] = v; if (v) bai.voffs.push_back(v); }
        p += 8 * (size_t)n_intv;
    }
    std::sort(bai.voffs.begin(), bai.voffs.end());
    bai.voffs.erase(std::unique(bai.voffs.begin(), bai.voffs.end()), bai.voffs.end());
    return 0;
}

struct VRange { uint64_t vb, ve; };                        // records from virtual offset vb up to ve (both record aligned)

// Which parts of the file can hold a read overlapping [beg, end) of reference tid: the chunks of every bin that overlaps the
// region, clipped below by the linear index entry of the region's first 16 kb tile -- hts_itr_query's rule (hts.c:3049-3090;
// reg2bins hts.c:2920, the linear-index lower bound hts.c:3099-3111).  A superset of what the iterator yields: it still
// filters record by record (hts.c:3913-3916), and so does the device search.
void bai_query(const BaiRefIdx &R, long long beg, long long end, std::vector<VRange> &out) {
    if (end <= beg || R.bins.empty()) return;
    if (beg < 0) beg = 0;
    if (end > (1ll << 29)) end = 1ll << 29;                 // the BAI binning scheme covers 2^29 bases
    if (beg >= end) return;
    uint64_t min_off = 0;
    if (!R.lin.empty()) {
        const size_t t = (size_t)(beg >> 14);
        min_off = t < R.lin.size() ? R.lin[t] : R.lin.back();
        if (min_off == 0 && t < R.lin.size()) for (size_t k = t; k-- > 0;) if (R.lin[k]) { min_off = R.lin[k]; break; }   // hts.c: walk back over empty tiles
    }
    const long long e1 = end - 1;
    int t = 0;
    for (int level = 0; level <= 5; ++level) {
        const int shift = 29 - 3 * level;                   // 26, 23, 20, 17, 14 for levels 1..5; level 0: the whole range
        const uint32_t b0 = (uint32_t)t + (uint32_t)(level ? (beg >> shift) : 0), b1 = (uint32_t)t + (uint32_t)(level ? (e1 >> shift) : 0);
        auto it = std::lower_bound(R.bins.begin(), R.bins.end(), b0, [](const BaiBin &a, uint32_t v) { return a.bin < v; });
        for (; it != R.bins.end() && it->bin <= b1; ++it)
            for (uint32_t k = 0; k < it->count; ++k) {
                const uint64_t vb = R.chunks[2 * (size_t)(it->first + k)], ve = R.chunks[2 * (size_t)(it->first + k) + 1];
                if (ve > min_off && ve > vb) out.push_back(VRange{std::max(vb, min_off), ve});
            }
        t += 1 << (3 * level);
    }
}

// sort, then merge ranges that overlap, touch, or lie within one BGZF block of each other (the records in between come along:
// extra reads are harmless, the per-window selection happens on the device)
void merge_ranges(std::vector<VRange> &r) {
    std::sort(r.begin(), r.end(), [](const VRange &a, const VRange &b) { return a.vb < b.vb; });
    size_t w = 0;
    for (size_t i = 0; i < r.size(); ++i) {
        if (w && (r[i].vb >> 16) <= (r[w - 1].ve >> 16) + 0x10000ull) r[w - 1].ve = std::max(r[w - 1].ve, r[i].ve);
        else r[w++] = r[i];
    }
    r.resize(w);
}

double ms_since(std::chrono::steady_clock::time_point t0) {
    return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
}

struct Span { size_t fb, fe; };                            // bytes [fb, fe) of the file: whole BGZF blocks

}  // namespace

// The loader proper.  ranges == nullptr: every record of the file.  Otherwise: the records of the given (sorted, merged) ranges.
//
// Pipeline (one pass over the needed bytes of the file):
//   host   T copier threads pread() the next chunk (<= 32 MB) of the file into a pinned ring buffer;  this thread then walks the
//          chunk's BGZF headers / footers (block table: payload offset, ISIZE -> each block's slice of the inflated image),
//   DMA    ships the chunk and its block descriptors on the ctx stream,
//   GPU    k_bgzf_inflate for the blocks that are complete so far, on one of a few side streams that waits for the chunk's event:
//          chunk k inflates while chunk k+1 uploads and chunk k+2 is read.
// then CRC, record chain, extraction as before.
static int load_bam_ranges(svtk_ctx *ctx, const char *bam_path, const char *bai_path, const Bai *bai_in, const std::vector<VRange> *ranges,
                           svtk_bam_stats *stats) {
    Ctx *c = reinterpret_cast<Ctx *>(ctx);
    if (stats) memset(stats, 0, sizeof(*stats));
    if (int rc = set_device(c)) return rc;
    const auto t_all = std::chrono::steady_clock::now();
    c->have_reads = false;
    const bool whole = ranges == nullptr;

    struct Fd { int fd = -1; ~Fd() { if (fd >= 0) ::close(fd); } } file;
    file.fd = ::open(bam_path, O_RDONLY);
    if (file.fd < 0) return fail(SVTK_ERR_IO, "bam: cannot open %s", bam_path);
    struct stat fst;
    if (fstat(file.fd, &fst) != 0) return fail(SVTK_ERR_IO, "bam: cannot stat %s", bam_path);
    const size_t fsize = (size_t)fst.st_size;
    if (fsize == 0) return fail(SVTK_ERR_IO, "bam: %s is empty", bam_path);

    Bai bai_local;
    const Bai *bai = bai_in;
    if (!bai) {
        std::string bp = bai_path ? std::string(bai_path) : std::string(bam_path) + ".bai";
        if (!bai_path && access(bp.c_str(), R_OK) != 0) {
            std::string alt(bam_path);
            if (alt.size() > 4 && alt.compare(alt.size() - 4, 4, ".bam") == 0) { alt.replace(alt.size() - 4, 4, ".bai"); if (access(alt.c_str(), R_OK) == 0) bp = alt; }
        }
        if (int rc = read_bai(bp.c_str(), bai_local)) return rc;
        bai = &bai_local;
    }

    // ---- header (host): the first bytes of the file
    unsigned long long hdr_end = 0;
    {
        size_t take = std::min<size_t>(fsize, 1u << 20);
        for (;;) {
            std::vector<uint8_t> hb(take);
            if (!pread_all(file.fd, hb.data(), take, 0)) return fail(SVTK_ERR_IO, "bam: cannot read %s", bam_path);
            bool need_more = false;
            const int rc = read_header(hb.data(), take, take == fsize, c->bam_ref_names, c->bam_ref_lens, &hdr_end, &need_more);
            if (rc == 0) break;
            if (!need_more || take == fsize) return rc;
            take = std::min<size_t>(fsize, take * 8);
        }
    }
    const int n_ref = (int)c->bam_ref_names.size();
    if (n_ref != (int)bai->refs.size()) return fail(SVTK_ERR_IO, "bam: index has %zu references, BAM header %d", bai->refs.size(), n_ref);

    // ---- which bytes: whole BGZF blocks around every range
    std::vector<Span> spans;
    if (whole) spans.push_back(Span{0, fsize});
    else
        for (const VRange &r : *ranges) {
            const size_t fb = (size_t)(r.vb >> 16);
            size_t fe = (size_t)(r.ve >> 16);
            if (fb >= fsize || fe > fsize) return fail(SVTK_ERR_IO, "bam: index offset beyond the end of the file (stale .bai?)");
            if ((r.ve & 0xffffu) != 0) {                            // the range ends inside a block: that block comes along whole
                uint8_t h[18 + 256];
                const size_t av = std::min<size_t>(sizeof h, fsize - fe);
                uint32_t bsize = 0, xlen = 0;
                if (!pread_all(file.fd, h, av, fe)) return fail(SVTK_ERR_IO, "bam: cannot read %s", bam_path);
                if (parse_bgzf_header(h, av, fe, &bsize, &xlen)) return fail(SVTK_ERR_IO, "bam: index offset %zu does not point at a BGZF block (stale .bai?)", fe);
                fe = std::min(fsize, fe + bsize);
            }
            if (!spans.empty() && fb <= spans.back().fe) spans.back().fe = std::max(spans.back().fe, fe);
            else if (fe > fb) spans.push_back(Span{fb, fe});
        }
    size_t image_bytes = 0;
    for (const Span &sp : spans) image_bytes += sp.fe - sp.fb;
    if (image_bytes >= (1ull << 34) - (1u << 20))
        return fail(SVTK_ERR_ARG, "bam: %zu bytes to load at once; 16 GiB and more must be split by region (svtk_load_bam_refs / svtk_load_bam_windows; "
                                  "the decoder indexes the shipped bytes with 32-bit words)", image_bytes);

    // ---- device + pinned buffers
    const size_t CH = std::min<size_t>(32u << 20, std::max<size_t>(1u << 20, (image_bytes + 3) / 4 + 65536));    // small files: small pinned buffers
    constexpr int NSTAGE = 3, NSIDE = 4, NCHEV = 8;
    for (int i = 0; i < NSTAGE; ++i) {
        if (int rc = c->bam_stage[i].reserve(CH)) return rc;
        if (!c->bam_stage_ev[i]) CU(cudaEventCreateWithFlags(&c->bam_stage_ev[i], cudaEventDisableTiming));
    }
    for (int i = 0; i < NSIDE; ++i) {
        if (!c->bam_side[i]) CU(cudaStreamCreateWithFlags(&c->bam_side[i], cudaStreamNonBlocking));
        if (!c->bam_side_ev[i]) CU(cudaEventCreateWithFlags(&c->bam_side_ev[i], cudaEventDisableTiming));
    }
    for (int i = 0; i < NCHEV; ++i) if (!c->bam_chunk_ev[i]) CU(cudaEventCreateWithFlags(&c->bam_chunk_ev[i], cudaEventDisableTiming));
    for (int i = 0; i < 5; ++i) if (!c->bam_ev[i]) CU(cudaEventCreate(&c->bam_ev[i]));
    if (int rc = c->bam_comp.reserve(image_bytes + BZ_OVERREAD + 64)) return rc;
    // capacity guesses for what is only known after the scan; a file that outgrows them takes the unpipelined route below
    size_t blk_cap = std::max<size_t>(c->bam_blocks.cap / sizeof(BgzfBlock), image_bytes / 8192 + 4096);
    if (int rc = c->bam_blocks.reserve(sizeof(BgzfBlock) * blk_cap)) return rc;
    if (int rc = c->bam_tab.reserve(sizeof(BgzfBlock) * blk_cap)) return rc;
    size_t infl_cap = std::max<size_t>(c->bam_infl.cap, 2 * image_bytes + (16u << 20));
    if (int rc = c->bam_infl.reserve(infl_cap)) return rc;
    infl_cap = c->bam_infl.cap - 128;
    constexpr int MAX_LAUNCH = 4096;
    if (int rc = c->bam_ctl.reserve(256 + 4 * MAX_LAUNCH)) return rc;
    if (int rc = c->bam_h.reserve(256 + 8 * (size_t)(n_ref + 1))) return rc;
    CU(cudaMemsetAsync(c->bam_ctl.p, 0, 256 + 4 * MAX_LAUNCH, c->stream));
    CU(cudaMemsetAsync((char *)c->bam_comp.p + (image_bytes & ~(size_t)3), 0, BZ_OVERREAD + 64 - (image_bytes & 3), c->stream));   // the decoder looks a few bytes past the end
    unsigned int *ctl = (unsigned int *)c->bam_ctl.p;
    int sms = 148;
    CU(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, c->device));
    static unsigned long long attr_seen = 0;
    if (first_use_on_device(attr_seen)) CU(cudaFuncSetAttribute(k_bgzf_inflate, cudaFuncAttributeMaxDynamicSharedMemorySize, (32 * BZ_TABLE_WORDS + BZ_LUT_WORDS) * 4));

    // lanes per warp of an inflate launch (see k_bgzf_inflate): thin warps finish a block sooner, wide warps fit more blocks on
    // the machine.  A launch should fit the 20 resident single-warp CTAs per SM the kernel's 96 registers allow.
    long long want_warps = (long long)sms * 20;
    if (const char *e = getenv("SVTK_BZ_WARPS")) if (*e) want_warps = std::max(1ll, atoll(e));
    int forced_lanes = 0;
    if (const char *e = getenv("SVTK_BZ_LANES")) if (*e) { const int v = atoi(e); if (v >= 1 && v <= 32 && (v & (v - 1)) == 0) forced_lanes = v; }
    auto lanes_for = [&](size_t blocks) {
        int l = 1;
        while (l < 32 && (long long)((blocks + l - 1) / l) > want_warps) l <<= 1;
        return forced_lanes ? forced_lanes : l;
    };
    int lanes = lanes_for(image_bytes / 40000 + 1);
    // Upload and inflate overlap in a FEW big pieces: a launch lasts as long as its slowest block (10 - 30 ms whatever its size:
    // 57 launches of 32 MB each took 450 ms against 71 ms for one launch of everything, profiles/r02a_bam_sweep.jsonl), so a
    // piece must upload for about as long as it inflates.  Files below 768 MB: one launch after the upload.
    size_t pipe_bytes = std::max<size_t>(384u << 20, image_bytes / 3 + 1);     // B200, 1.9 GB file: 4 pieces 98 ms, 3 pieces 88 ms, 1 piece 126 ms
    if (const char *e = getenv("SVTK_BAM_PIPE_MB")) if (*e) pipe_bytes = (size_t)std::max(1ll, atoll(e)) << 20;
    size_t first_bytes = pipe_bytes;                                            // the first piece may be smaller: the GPU starts earlier
    if (const char *e = getenv("SVTK_BAM_PIPE_FIRST_MB")) if (*e) first_bytes = (size_t)std::max(1ll, atoll(e)) << 20;
    bool pipelined = image_bytes >= (768u << 20);
    if (const char *e = getenv("SVTK_BAM_PIPE")) if (*e) pipelined = (*e == '1');
    if (pipelined) {      // concurrent launches need their own token slices: room for the whole load, by the capacity guesses
        if (int rc = c->bam_tok.reserve(8 * (size_t)(bz_tok_capacity(infl_cap, blk_cap) + 4))) return rc;
        if (int rc = c->bam_ntok.reserve(4 * std::max<size_t>(blk_cap, 1))) return rc;
    }
    size_t last_launch_off = 0;

    std::vector<uint64_t> coff, cend;                      // file offset of every loaded block (sorted), and of its end
    std::vector<BgzfBlock> tab_overflow;                   // descriptors beyond the pinned table's first guess
    BgzfBlock *tab = (BgzfBlock *)c->bam_tab.p;            // pinned: descriptors are DMA'd from here
    size_t n_blk = 0, launched = 0, shipped_tab = 0;
    auto tab_at = [&](size_t i) -> const BgzfBlock & { return tab[i]; };
    unsigned long long total = 0;
    int n_launch = 0;
    bool first_launch = true;
    // both phases for blocks [b0, b1): in waves of <= ~3 GB of inflated bytes, so that the match-token buffer (8 bytes per 3
    // inflated bytes at worst) stays bounded whatever the file size
    auto launch_inflate = [&](size_t b0, size_t b1, cudaStream_t st, bool whole_buffer) -> int {
        if (!whole_buffer) if (int rc = c->bam_ntok.reserve(4 * std::max<size_t>(std::max(n_blk, blk_cap), 1))) return rc;
        const unsigned long long WAVE = whole_buffer ? ~0ull : (3ull << 30);
        while (b0 < b1) {
            if (n_launch >= MAX_LAUNCH) return fail(SVTK_ERR_STATE, "bam: too many inflate launches");
            size_t be = b0 + 1;
            while (be < b1 && tab_at(be).out_off + tab_at(be).isize - tab_at(b0).out_off <= WAVE) ++be;
            const size_t nb = be - b0;
            const unsigned long long wave_bytes = tab_at(be - 1).out_off + tab_at(be - 1).isize - tab_at(b0).out_off;
            // whole_buffer: the token buffer was sized for the whole load up front (launches run side by side, each in its own slice)
            const unsigned long long tok_sub = whole_buffer ? 0ull : bz_tok_offset(tab_at(b0).out_off, b0);
            if (!whole_buffer) if (int rc = c->bam_tok.reserve(8 * (size_t)(bz_tok_capacity(wave_bytes, nb) + 4))) return rc;
            lanes = lanes_for(nb);
            const int smem = (lanes * BZ_TABLE_WORDS + BZ_LUT_WORDS) * 4;
            int per_sm = 1;
            CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_bgzf_inflate, 32, smem));
            const unsigned grid = (unsigned)std::min<size_t>((size_t)sms * std::max(per_sm, 1), (nb + lanes - 1) / lanes);
            k_bgzf_inflate<<<grid, 32, smem, st>>>((const uint32_t *)c->bam_comp.p, (const BgzfBlock *)c->bam_blocks.p, (unsigned)b0, (unsigned)be,
                                                   (uint8_t *)c->bam_infl.p, (unsigned long long *)c->bam_tok.p, tok_sub, (uint32_t *)c->bam_ntok.p, ctl,
                                                   ctl + 64 + n_launch, lanes);
            CU(cudaGetLastError());
            const unsigned rgrid = (unsigned)std::min<size_t>((size_t)sms * 8, (nb + 7) / 8);
            k_bgzf_resolve<<<rgrid, 256, 0, st>>>((const BgzfBlock *)c->bam_blocks.p, (unsigned)b0, (unsigned)be, (uint8_t *)c->bam_infl.p,
                                                  (const unsigned long long *)c->bam_tok.p, tok_sub, (const uint32_t *)c->bam_ntok.p);
            CU(cudaGetLastError());
            ++n_launch;
            c->launches += 2;
            b0 = be;
        }
        return 0;
    };

    // ---- the pipeline
    const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
    int T = (int)std::min(16u, std::max(1u, hw * 3 / 4));        // pread threads (the calling thread issues the copies and walks the block headers)
    if (const char *e = getenv("SVTK_BAM_THREADS")) if (*e) T = std::max(1, std::min(64, atoi(e)));
    bool fits = true;                                       // the capacity guesses still hold
    CU(cudaEventRecord(c->bam_ev[0], c->stream));
    size_t img_off = 0;                                     // bytes of the image shipped so far
    int k = 0;                                              // chunk counter
    // scanner state: the block whose header comes next
    size_t scan_span = 0, scan_off = spans.empty() ? 0 : spans[0].fb;
    size_t span_img0 = 0;                                   // image offset of the current scan span's first byte
    uint8_t carry[65536 + 64];                              // head of a block that straddles two chunks
    // reader pool: T threads pread() their slice of every chunk into the pinned ring, up to NSTAGE chunks ahead of the DMA
    struct Piece { size_t si, file_off, n; };
    std::vector<Piece> pieces;
    for (size_t si = 0; si < spans.size(); ++si)
        for (size_t d = spans[si].fb; d < spans[si].fe; d += CH) pieces.push_back(Piece{si, d, std::min(CH, spans[si].fe - d)});
    struct Pool {
        std::vector<std::thread> th;
        std::atomic<long> released{0};
        std::atomic<int> failed{0};
        std::unique_ptr<std::atomic<int>[]> arrived;
        ~Pool() { failed.store(1); for (auto &x : th) if (x.joinable()) x.join(); }
    } pool;
    pool.arrived.reset(new std::atomic<int>[pieces.size() + 1]);
    for (size_t i = 0; i <= pieces.size(); ++i) pool.arrived[i].store(0);
    pool.released.store(std::min<long>((long)pieces.size(), NSTAGE));
    {
        const int fd = file.fd;
        char *ring[NSTAGE];
        for (int i = 0; i < NSTAGE; ++i) ring[i] = (char *)c->bam_stage[i].p;
        const Piece *pc = pieces.data();
        const long np = (long)pieces.size();
        for (int t = 0; t < T; ++t)
            pool.th.emplace_back([&pool, pc, np, fd, t, T, ring]() {
                for (long q = 0; q < np; ++q) {
                    while (pool.released.load(std::memory_order_acquire) <= q) { if (pool.failed.load()) return; std::this_thread::yield(); }
                    const size_t n = pc[q].n, per = ((n + (size_t)T - 1) / (size_t)T + 4095) & ~(size_t)4095, a = per * (size_t)t;
                    if (a < n && !pread_all(fd, ring[q % NSTAGE] + a, std::min(per, n - a), pc[q].file_off + a)) { pool.failed.store(1); return; }
                    pool.arrived[q].fetch_add(1, std::memory_order_release);
                }
            });
    }
    for (size_t pi = 0; pi < pieces.size(); ++pi) {
        {
            const size_t si = pieces[pi].si, done = pieces[pi].file_off, n = pieces[pi].n;
            const int sb = k % NSTAGE;
            char *st = (char *)c->bam_stage[sb].p;
            while (pool.arrived[pi].load(std::memory_order_acquire) < T) {
                if (pool.failed.load()) return fail(SVTK_ERR_IO, "bam: cannot read %s", bam_path);
                std::this_thread::yield();
            }
            CU(cudaMemcpyAsync((char *)c->bam_comp.p + img_off, st, n, cudaMemcpyHostToDevice, c->stream));
            CU(cudaEventRecord(c->bam_stage_ev[sb], c->stream));
            if (k >= 1) {                                             // the previous chunk's DMA is over (or nearly): its buffer takes chunk k - 1 + NSTAGE
                CU(cudaEventSynchronize(c->bam_stage_ev[(k - 1) % NSTAGE]));
                pool.released.store(std::min<long>((long)pieces.size(), (long)k - 1 + NSTAGE + 1), std::memory_order_release);
            }
            // ---- block table of the bytes staged so far (this chunk: file bytes [done, done + n) of span si)
            const size_t have_to = done + n;
            while (scan_span == si && scan_off < spans[si].fe) {
                // header + footer of the block at scan_off: from the staged chunk when it lies inside, else through a small read
                const uint8_t *h;
                size_t av;
                if (scan_off >= done && scan_off + 18 + 256 <= have_to) { h = (const uint8_t *)st + (scan_off - done); av = have_to - scan_off; }
                else {
                    av = std::min<size_t>(18 + 256, fsize - scan_off);
                    if (scan_off + 18 > have_to) break;                 // header not staged yet
                    if (!pread_all(file.fd, carry, av, scan_off)) return fail(SVTK_ERR_IO, "bam: cannot read %s", bam_path);
                    h = carry;
                }
                uint32_t bsize = 0, xlen = 0;
                if (int rc = parse_bgzf_header(h, std::min<size_t>(av, 65536), scan_off, &bsize, &xlen)) return rc;
                if (scan_off + bsize > spans[si].fe) return fail(SVTK_ERR_IO, "bam: bad BGZF block size at byte %zu", scan_off);
                if (scan_off + bsize > have_to) break;                  // the block ends in the next chunk
                uint32_t isize;
                if (scan_off >= done) isize = le32((const uint8_t *)st + (scan_off + bsize - 4 - done));
                else {                                                  // straddler: its tail is in this chunk
                    const size_t fo = scan_off + bsize - 4;
                    if (fo >= done) isize = le32((const uint8_t *)st + (fo - done));
                    else { uint8_t f4[4]; if (!pread_all(file.fd, f4, 4, fo)) return fail(SVTK_ERR_IO, "bam: cannot read %s", bam_path); isize = le32(f4); }
                }
                if (isize > 65536u) return fail(SVTK_ERR_IO, "bam: BGZF block at byte %zu claims %u inflated bytes", scan_off, isize);
                if (n_blk >= 0xfffffff0ull) return fail(SVTK_ERR_ARG, "bam: more than 2^32 BGZF blocks: split the load");
                if (fits && (n_blk >= blk_cap || total + isize > infl_cap)) fits = false;
                BgzfBlock b;
                b.in_beg = span_img0 + (scan_off - spans[si].fb) + 12 + xlen; b.in_len = bsize - xlen - 20; b.isize = isize; b.out_off = total;
                if (fits) tab[n_blk] = b; else tab_overflow.push_back(b);
                coff.push_back(scan_off); cend.push_back(scan_off + bsize);
                ++n_blk;
                total += isize;
                scan_off += bsize;
            }
            if (scan_span == si && scan_off >= spans[si].fe && si + 1 < spans.size()) {
                span_img0 += spans[si].fe - spans[si].fb;
                scan_span = si + 1; scan_off = spans[si + 1].fb;
            }
            img_off += n;
            // ---- ship the new descriptors, launch their inflate behind this chunk's bytes
            if (fits && pipelined && n_blk > launched && (img_off - last_launch_off >= (first_launch ? first_bytes : pipe_bytes) || img_off == image_bytes)) {
                last_launch_off = img_off;
                CU(cudaMemcpyAsync((BgzfBlock *)c->bam_blocks.p + shipped_tab, tab + shipped_tab, sizeof(BgzfBlock) * (n_blk - shipped_tab), cudaMemcpyHostToDevice, c->stream));
                shipped_tab = n_blk;
                cudaEvent_t ev = c->bam_chunk_ev[k % NCHEV];
                CU(cudaEventRecord(ev, c->stream));
                cudaStream_t side = c->bam_side[n_launch % NSIDE];
                CU(cudaStreamWaitEvent(side, ev, 0));
                if (first_launch) { CU(cudaEventRecord(c->bam_ev[1], side)); first_launch = false; }
                if (int rc = launch_inflate(launched, n_blk, side, true)) return rc;
                launched = n_blk;
            }
            ++k;
        }
    }
    if (scan_span + 1 < spans.size() || (!spans.empty() && scan_off != spans.back().fe))
        return fail(SVTK_ERR_IO, "bam: BGZF blocks do not tile the file (stopped at byte %zu)", scan_off);
    CU(cudaEventRecord(c->bam_ev[2], c->stream));           // end of the upload
    if (!fits) {
        // the guesses were too small: exact buffers now, the whole inflate after the upload (everything is on the device already)
        for (int i = 0; i < NSIDE; ++i) CU(cudaStreamSynchronize(c->bam_side[i]));
        CU(cudaStreamSynchronize(c->stream));
        std::vector<BgzfBlock> all(n_blk);
        const size_t in_tab = n_blk - tab_overflow.size();
        memcpy(all.data(), tab, sizeof(BgzfBlock) * in_tab);
        if (!tab_overflow.empty()) memcpy(all.data() + in_tab, tab_overflow.data(), sizeof(BgzfBlock) * tab_overflow.size());
        tab_overflow.clear();
        if (int rc = c->bam_blocks.reserve(sizeof(BgzfBlock) * n_blk)) return rc;
        if (int rc = c->bam_tab.reserve(sizeof(BgzfBlock) * n_blk)) return rc;
        if (int rc = c->bam_infl.reserve((size_t)total + 256)) return rc;
        tab = (BgzfBlock *)c->bam_tab.p;
        memcpy(tab, all.data(), sizeof(BgzfBlock) * n_blk);
        launched = 0; shipped_tab = 0;
        CU(cudaMemsetAsync(c->bam_ctl.p, 0, 256 + 4 * MAX_LAUNCH, c->stream));
        n_launch = 0;
    }
    if (n_blk > shipped_tab) {
        CU(cudaMemcpyAsync((BgzfBlock *)c->bam_blocks.p + shipped_tab, tab + shipped_tab, sizeof(BgzfBlock) * (n_blk - shipped_tab), cudaMemcpyHostToDevice, c->stream));
        shipped_tab = n_blk;
    }
    if (first_launch) { CU(cudaEventRecord(c->bam_ev[1], c->stream)); first_launch = false; }
    if (n_blk > launched) { if (int rc = launch_inflate(launched, n_blk, c->stream, pipelined && fits)) return rc; launched = n_blk; }
    for (int i = 0; i < NSIDE; ++i) {                       // join the side streams
        CU(cudaEventRecord(c->bam_side_ev[i], c->bam_side[i]));
        CU(cudaStreamWaitEvent(c->stream, c->bam_side_ev[i], 0));
    }
    CU(cudaMemsetAsync((char *)c->bam_infl.p + total, 0, 64, c->stream));     // the record kernels read whole words: zeros behind the last byte
    CU(cudaEventRecord(c->bam_ev[3], c->stream));           // end of the inflate
    if (stats) stats->lanes_per_warp = (uint32_t)lanes;
    const uint32_t *infl = (const uint32_t *)c->bam_infl.p;
    if (n_blk) {
        const unsigned grid = (unsigned)std::min<size_t>((size_t)sms * 8, (n_blk + 7) / 8);
        k_bgzf_crc<<<grid, 256, 0, c->stream>>>((const uint32_t *)c->bam_comp.p, (const BgzfBlock *)c->bam_blocks.p, (unsigned)n_blk, (const uint8_t *)c->bam_infl.p, ctl);
        CU(cudaGetLastError());
        c->launches += 1;
    }

    // ---- pieces of the record chain: every range, cut at the record-aligned offsets the index holds
    auto to_linear = [&](uint64_t v, unsigned long long *lin) -> int {
        const uint64_t co = v >> 16, uo = v & 0xffffu;
        auto it = std::lower_bound(coff.begin(), coff.end(), co);
        if (it == coff.end() || *it != co) {
            // just past the last loaded block of a span: the end of the previous block
            if (uo == 0 && it != coff.begin() && cend[(size_t)(it - coff.begin()) - 1] == co) {
                const BgzfBlock &pb = tab[(size_t)(it - coff.begin()) - 1];
                *lin = pb.out_off + pb.isize;
                return 0;
            }
            return fail(SVTK_ERR_IO, "bam: index offset %llu does not point at a loaded BGZF block (stale .bai?)", (unsigned long long)co);
        }
        const BgzfBlock &b = tab[(size_t)(it - coff.begin())];
        if (uo > b.isize) return fail(SVTK_ERR_IO, "bam: index offset beyond its block (stale .bai?)");
        *lin = b.out_off + uo;
        return 0;
    };
    std::vector<unsigned long long> splits;                 // (begin, end) pairs, linear offsets into the inflated image
    auto add_range = [&](unsigned long long lb, unsigned long long le, uint64_t vb, uint64_t ve) -> int {
        if (lb >= le) return 0;
        unsigned long long prev = lb;
        auto it = std::upper_bound(bai->voffs.begin(), bai->voffs.end(), vb);
        for (; it != bai->voffs.end() && *it < ve; ++it) {
            unsigned long long lin = 0;
            if (int rc = to_linear(*it, &lin)) return rc;
            if (lin > prev && lin < le) { splits.push_back(prev); splits.push_back(lin); prev = lin; }
        }
        splits.push_back(prev); splits.push_back(le);
        return 0;
    };
    if (whole) {
        if (hdr_end > total) return fail(SVTK_ERR_IO, "bam: header / index offsets beyond the file");
        // virtual offset of the first record: the block holding inflated byte hdr_end
        size_t j = 0;
        while (j + 1 < n_blk && tab[j + 1].out_off <= hdr_end) ++j;
        const uint64_t v_first = n_blk ? (coff[j] << 16 | (hdr_end - tab[j].out_off)) : 0;
        if (int rc = add_range(hdr_end, total, v_first, ~0ull)) return rc;
    } else
        for (const VRange &r : *ranges) {
            unsigned long long lb = 0, le = 0;
            if (int rc = to_linear(r.vb, &lb)) return rc;
            if (int rc = to_linear(r.ve, &le)) return rc;
            if (lb > le || le > total) return fail(SVTK_ERR_IO, "bam: index offsets beyond the file (stale .bai?)");
            if (int rc = add_range(lb, le, r.vb, r.ve)) return rc;
        }
    const int n_seg = (int)(splits.size() / 2);
    if (int rc = c->bam_splits.reserve(8 * std::max<size_t>(splits.size(), 2))) return rc;
    if (int rc = c->bam_seg.reserve(16 * (size_t)std::max(n_seg, 1))) return rc;
    if (int rc = c->bam_h.reserve(256 + 8 * (size_t)(n_ref + 1) + 16 * (size_t)std::max(n_seg, 1))) return rc;
    if (n_seg) if (int rc = upload(c, c->bam_splits.p, splits.data(), 8 * splits.size())) return rc;

    // ---- record chain: count, prefix (host: a few thousand pieces), fill
    unsigned int *hctl = (unsigned int *)c->bam_h.p;
    long long n_reads = 0, n_words = 0;
    std::vector<unsigned long long> base(2 * (size_t)std::max(n_seg, 1), 0ull);
    if (n_seg > 0) {
        k_bam_walk<false><<<(n_seg + 127) / 128, 128, 0, c->stream>>>(infl, (const unsigned long long *)c->bam_splits.p, n_seg, (ulonglong2 *)c->bam_seg.p,
                                                                      nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, ctl);
        CU(cudaGetLastError());
        c->launches += 1;
    }
    CU(cudaMemcpyAsync(hctl, ctl, 64, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    if (hctl[BAM_CTL_ZERR]) return fail(SVTK_ERR_IO, "bam: BGZF block %u does not inflate (code %u)", hctl[BAM_CTL_ZERR] - 1, hctl[BAM_CTL_ZCODE]);
    if (hctl[BAM_CTL_CRC]) return fail(SVTK_ERR_IO, "bam: CRC32 checksum mismatch in BGZF block %u", hctl[BAM_CTL_CRC] - 1);
    if (hctl[BAM_CTL_WERR]) return fail(SVTK_ERR_IO, "bam: record chain broken in piece %u of %d (code %u): corrupt BAM or stale index", hctl[BAM_CTL_WSEG], n_seg, hctl[BAM_CTL_WERR]);
    if (n_seg > 0) {
        CU(cudaMemcpyAsync(c->bam_h.p, c->bam_seg.p, 16 * (size_t)n_seg, cudaMemcpyDeviceToHost, c->stream));
        CU(cudaStreamSynchronize(c->stream));
        const unsigned long long *cnt = (const unsigned long long *)c->bam_h.p;
        for (int s = 0; s < n_seg; ++s) { base[2 * s] = (unsigned long long)n_reads; base[2 * s + 1] = (unsigned long long)n_words; n_reads += (long long)cnt[2 * s]; n_words += (long long)cnt[2 * s + 1]; }
    }
    if (n_reads >= (1ll << 31) - 1024) return fail(SVTK_ERR_ARG, "bam: more than 2^31 reads per device: split the load");
    const size_t nr = (size_t)n_reads, nw = (size_t)n_words;
    if (int rc = c->r_pos.reserve(4 * std::max<size_t>(nr, 1))) return rc;
    if (int rc = c->r_end.reserve(4 * std::max<size_t>(nr, 1))) return rc;
    if (int rc = c->r_ncig.reserve(4 * std::max<size_t>(nr, 1))) return rc;
    if (int rc = c->r_off.reserve(8 * std::max<size_t>(nr, 1))) return rc;
    if (int rc = c->r_cigar.reserve(4 * nw + 16)) return rc;
    if (int rc = c->bam_rec.reserve(28 * std::max<size_t>(nr, 1))) return rc;     // per read: record offset, CIGAR source, SEQ source (u64 each), l_seq (u32)
    if (int rc = c->tmp_a.reserve(4 * std::max<size_t>(nr, 1))) return rc;        // tid per read
    if (int rc = c->tmp_b.reserve(8 * (size_t)(n_ref + 1))) return rc;
    unsigned long long *rec_off = (unsigned long long *)c->bam_rec.p, *cig_src = rec_off + nr, *seq_src = cig_src + nr;
    uint32_t *l_seq = (uint32_t *)(seq_src + nr);
    if (n_reads > 0) {
        if (int rc = upload(c, c->bam_seg.p, base.data(), 16 * (size_t)n_seg)) return rc;
        k_bam_walk<true><<<(n_seg + 127) / 128, 128, 0, c->stream>>>(infl, (const unsigned long long *)c->bam_splits.p, n_seg, (ulonglong2 *)c->bam_seg.p, rec_off,
                                                                     cig_src, (uint32_t *)c->r_ncig.p, (unsigned long long *)c->r_off.p, seq_src, l_seq, ctl);
        CU(cudaGetLastError());
        const long long threads = n_reads * 32;
        k_bam_extract<<<(unsigned)((threads + 255) / 256), 256, 0, c->stream>>>(infl, rec_off, cig_src, (const uint32_t *)c->r_ncig.p,
                                                                                (const unsigned long long *)c->r_off.p, n_reads, (uint32_t *)c->r_pos.p,
                                                                                (uint32_t *)c->r_end.p, (int32_t *)c->tmp_a.p, (uint32_t *)c->r_cigar.p);
        CU(cudaGetLastError());
        c->launches += 2;
    }
    k_bam_tid_off<<<(unsigned)((n_reads + 1 + 255) / 256), 256, 0, c->stream>>>((const int32_t *)c->tmp_a.p, (const uint32_t *)c->r_pos.p, n_reads, n_ref,
                                                                               (long long *)c->tmp_b.p, ctl);
    CU(cudaGetLastError());
    c->launches += 1;
    cudaEvent_t ev_done = c->bam_ev[4];
    CU(cudaEventRecord(ev_done, c->stream));
    std::vector<int64_t> tid_off((size_t)n_ref + 1, 0);
    CU(cudaMemcpyAsync(hctl, ctl, 64, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaMemcpyAsync(hctl + 16, c->tmp_b.p, 8 * (size_t)(n_ref + 1), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    if (hctl[BAM_CTL_ORDER]) return fail(SVTK_ERR_IO, "bam: records are not in coordinate order (or a refID is outside the header's dictionary)");
    memcpy(tid_off.data(), hctl + 16, 8 * (size_t)(n_ref + 1));
    c->bam_infl_bytes = total; c->bam_n_rec = n_reads;
    c->h2d_bytes += image_bytes + sizeof(BgzfBlock) * n_blk + 8 * splits.size();

    svtk_reads d{};
    d.n_tid = n_ref; d.n_reads = n_reads; d.n_words = n_words; d.tid_off = tid_off.data();
    d.pos = (const uint32_t *)c->r_pos.p; d.end = (const uint32_t *)c->r_end.p; d.n_cigar = (const uint32_t *)c->r_ncig.p;
    d.cigar_off = (const uint64_t *)c->r_off.p; d.cigar = (const uint32_t *)c->r_cigar.p;
    if (int rc = adopt_device_reads(c, &d)) return rc;
    c->reads_from_bam = true;
    if (stats) {
        float up = 0, inf = 0, span = 0;
        cudaEventElapsedTime(&up, c->bam_ev[0], c->bam_ev[2]); cudaEventElapsedTime(&inf, c->bam_ev[1], c->bam_ev[3]); cudaEventElapsedTime(&span, c->bam_ev[0], c->bam_ev[3]);
        stats->file_bytes = fsize; stats->image_bytes = image_bytes; stats->n_blocks = n_blk; stats->inflated_bytes = total; stats->n_reads = (uint64_t)n_reads;
        stats->n_words = (uint64_t)n_words; stats->n_ref = n_ref; stats->n_pieces = (uint32_t)n_seg; stats->host_scan_ms = 0.f;
        stats->upload_ms = up; stats->inflate_ms = inf; stats->load_span_ms = span; stats->n_spans = (uint32_t)spans.size(); stats->inflate_launches = (uint32_t)n_launch;
        float par = 0;
        cudaEventElapsedTime(&par, c->bam_ev[3], ev_done);
        stats->parse_ms = par; stats->total_ms = (float)ms_since(t_all);
    }
    return 0;
}

extern "C" int svtk_load_bam(svtk_ctx *ctx, const char *bam_path, const char *bai_path, svtk_bam_stats *stats) {
    if (!ctx || !bam_path) return fail(SVTK_ERR_ARG, "null argument");
    return load_bam_ranges(ctx, bam_path, bai_path, nullptr, nullptr, stats);
}

static int open_bai(const char *bam_path, const char *bai_path, Bai &bai) {
    std::string bp = bai_path ? std::string(bai_path) : std::string(bam_path) + ".bai";
    if (!bai_path && access(bp.c_str(), R_OK) != 0) {
        std::string alt(bam_path);
        if (alt.size() > 4 && alt.compare(alt.size() - 4, 4, ".bam") == 0) { alt.replace(alt.size() - 4, 4, ".bai"); if (access(alt.c_str(), R_OK) == 0) bp = alt; }
    }
    return read_bai(bp.c_str(), bai);
}

extern "C" int svtk_load_bam_refs(svtk_ctx *ctx, const char *bam_path, const char *bai_path, int32_t tid_lo, int32_t tid_hi, svtk_bam_stats *stats) {
    if (!ctx || !bam_path) return fail(SVTK_ERR_ARG, "null argument");
    if (tid_lo < 0) return fail(SVTK_ERR_ARG, "bad reference range");
    Bai bai;
    if (int rc = open_bai(bam_path, bai_path, bai)) return rc;
    if (tid_hi < tid_lo || tid_hi >= (int)bai.refs.size()) return fail(SVTK_ERR_ARG, "bam: reference range [%d, %d] outside the index's %zu references", tid_lo, tid_hi, bai.refs.size());
    std::vector<VRange> r;
    uint64_t vb = ~0ull, ve = 0;
    for (int t = tid_lo; t <= tid_hi; ++t) { vb = std::min(vb, bai.refs[(size_t)t].beg); ve = std::max(ve, bai.refs[(size_t)t].end); }
    if (vb < ve) r.push_back(VRange{vb, ve});
    return load_bam_ranges(ctx, bam_path, bai_path, &bai, &r, stats);
}

// Only what the given windows can see: per window the region sam_itr_queryi(idx, chrom - 1, start - 1, end - 1) would query
// (refinement.c:114), widened to the bin chunks that can hold an overlapping read.  A region shard of a single-contig job, or a
// sparse VCF against a whole-genome BAM, ships and inflates only those parts of the file.
extern "C" int svtk_load_bam_windows(svtk_ctx *ctx, const char *bam_path, const char *bai_path, const svtk_windows *w, svtk_bam_stats *stats) {
    if (!ctx || !bam_path || !w || w->n_win < 0) return fail(SVTK_ERR_ARG, "null argument");
    if (w->n_win > 0 && (!w->chrom || !w->start || !w->end)) return fail(SVTK_ERR_ARG, "null window array");
    Bai bai;
    if (int rc = open_bai(bam_path, bai_path, bai)) return rc;
    // union of the windows' regions per reference first (windows overlap heavily), then one index query per merged region
    std::vector<std::vector<std::pair<long long, long long>>> reg(bai.refs.size());
    for (int i = 0; i < w->n_win; ++i) {
        const long long tid = (long long)w->chrom[i] - 1;
        if (tid < 0 || tid >= (long long)bai.refs.size()) continue;
        const long long beg = (long long)(uint32_t)(w->start[i] - 1u), end = (long long)(uint32_t)(w->end[i] - 1u);   // uint32 arithmetic, then widened
        if (end < beg) continue;                               // hts_itr_query: NULL iterator
        reg[(size_t)tid].push_back({beg, std::max(end, beg + 1)});
    }
    std::vector<VRange> r;
    for (size_t t = 0; t < reg.size(); ++t) {
        auto &v = reg[t];
        if (v.empty()) continue;
        std::sort(v.begin(), v.end());
        long long b = v[0].first, e = v[0].second;
        for (size_t i = 1; i <= v.size(); ++i) {
            if (i < v.size() && v[i].first <= e + 16384) { e = std::max(e, v[i].second); continue; }
            bai_query(bai.refs[t], b, e, r);
            if (i < v.size()) { b = v[i].first; e = v[i].second; }
        }
    }
    merge_ranges(r);
    return load_bam_ranges(ctx, bam_path, bai_path, &bai, &r, stats);
}

// compressed bytes of the file under each window, by the .bai's linear index (16 kb tiles): what loading + inflating the reads of
// that window costs -- the work estimate region shards are balanced by (depth shows up as more bytes per base)
namespace svtk {
int bam_window_weights(const char *bam_path, const char *bai_path, int n, const int32_t *chrom, const uint32_t *start, const uint32_t *end, double *weight) {
    Bai bai;
    if (int rc = open_bai(bam_path, bai_path, bai)) return rc;
    for (int i = 0; i < n; ++i) {
        weight[i] = 1.0;
        const long long tid = (long long)chrom[i] - 1;
        if (tid < 0 || tid >= (long long)bai.refs.size()) continue;
        const BaiRefIdx &R = bai.refs[(size_t)tid];
        const long long beg = (long long)(uint32_t)(start[i] - 1u), e = (long long)(uint32_t)(end[i] - 1u);
        if (e < beg || R.lin.empty()) continue;
        const size_t t0 = std::min<size_t>((size_t)(beg >> 14), R.lin.size() - 1), t1 = (size_t)(e >> 14) + 1;
        uint64_t vb = R.lin[t0], ve = t1 < R.lin.size() ? R.lin[t1] : R.end;
        if (vb == 0) vb = R.beg;
        if (ve < vb) ve = vb;
        weight[i] = 65536.0 + (double)((ve >> 16) - (vb >> 16));
    }
    return 0;
}
}  // namespace svtk

extern "C" int svtk_bam_n_ref(const svtk_ctx *ctx) { return ctx ? (int)reinterpret_cast<const Ctx *>(ctx)->bam_ref_names.size() : 0; }

extern "C" int svtk_bam_ref(const svtk_ctx *ctx, int tid, const char **name, uint32_t *len) {
    const Ctx *c = reinterpret_cast<const Ctx *>(ctx);
    if (!c || tid < 0 || tid >= (int)c->bam_ref_names.size()) return fail(SVTK_ERR_ARG, "bad reference id");
    if (name) *name = c->bam_ref_names[(size_t)tid].c_str();
    if (len) *len = c->bam_ref_lens[(size_t)tid];
    return 0;
}

// the resident read set back on the host (any array may be NULL: call once for the sizes, then with buffers)
extern "C" int svtk_copy_reads(svtk_ctx *ctx, int64_t *n_tid, int64_t *n_reads, int64_t *n_words, int64_t *tid_off, uint32_t *pos, uint32_t *end,
                               uint32_t *n_cigar, uint64_t *cigar_off, uint32_t *cigar) {
    Ctx *c = reinterpret_cast<Ctx *>(ctx);
    if (!c) return fail(SVTK_ERR_ARG, "null ctx");
    if (!c->have_reads) return fail(SVTK_ERR_STATE, "no reads loaded");
    if (int rc = set_device(c)) return rc;
    const LaunchReads &R = c->reads;
    if (n_tid) *n_tid = R.n_tid;
    if (n_reads) *n_reads = R.n_reads;
    if (n_words) *n_words = c->n_words;
    CU(cudaStreamSynchronize(c->stream));
    const size_t nr = (size_t)R.n_reads;
    if (tid_off) memcpy(tid_off, c->h_tid_off.data(), 8 * (size_t)(R.n_tid + 1));
    if (pos && nr) CU(cudaMemcpy(pos, R.pos, 4 * nr, cudaMemcpyDeviceToHost));
    if (end && nr) CU(cudaMemcpy(end, R.end, 4 * nr, cudaMemcpyDeviceToHost));
    if (n_cigar && nr) CU(cudaMemcpy(n_cigar, R.n_cigar, 4 * nr, cudaMemcpyDeviceToHost));
    if (cigar_off && nr) CU(cudaMemcpy(cigar_off, R.cigar_off, 8 * nr, cudaMemcpyDeviceToHost));
    if (cigar && c->n_words) CU(cudaMemcpy(cigar, R.cigar, 4 * (size_t)c->n_words, cudaMemcpyDeviceToHost));
    return 0;
}
