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
#include <chrono>
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
__global__ void __launch_bounds__(32, SVTK_BZ_MINB) k_bgzf_inflate(const uint32_t *__restrict__ words, const BgzfBlock *__restrict__ blk, unsigned n_blocks,
                                                     uint8_t *__restrict__ out, unsigned int *__restrict__ ctl, int lanes) {
    extern __shared__ uint32_t bz_sm[];
    BzTables T;
    T.lit = reinterpret_cast<uint16_t *>(bz_sm + (threadIdx.x < (unsigned)lanes ? threadIdx.x : 0u) * BZ_TABLE_WORDS);
    T.dist = T.lit + (1 << BZ_LBITS);
    BzLane z;
    z.state = BZ_S_IDLE; z.err = BZ_OK;
    unsigned cur = 0xffffffffu;
    bool drained = threadIdx.x >= (unsigned)lanes;
    for (;;) {
        if (z.state == BZ_S_IDLE && !drained) {
            if (cur != 0xffffffffu && z.err) { if (atomicCAS(&ctl[BAM_CTL_ZERR], 0u, cur + 1u) == 0u) ctl[BAM_CTL_ZCODE] = (unsigned)z.err; }
            cur = atomicAdd(&ctl[BAM_CTL_NEXT], 1u);
            if (cur < n_blocks) {
                const BgzfBlock b = blk[cur];
                bz_begin(z, words, b.in_beg, b.in_beg + b.in_len, out + b.out_off, b.isize);
            } else { drained = true; z.err = BZ_OK; }
        }
        if (__all_sync(FULLMASK, drained)) break;
        // every lane passes here once per round: block switches (header + table build) on the lanes that need one, then a
        // burst of symbol steps that all lanes take together
        if (z.state == BZ_S_HEADER) bz_header(z, T);
#pragma unroll 1
        for (int it = 0; it < BZ_BURST; ++it)
            if (z.state == BZ_S_SYMBOLS) bz_step(z, T);
        bz_check_input(z);
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

// One thread per piece [splits[s], splits[s+1]) of the record chain.  FILL == false: seg[s] = {reads, CIGAR words}.
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
    unsigned long long off = splits[s];
    const unsigned long long end = splits[s + 1];
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

struct Mapped {                      // read-only mmap of a whole file
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
        madvise(m, n, MADV_SEQUENTIAL);
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

// BGZF block table (SAMv1 4.1): gzip member with the BC extra subfield carrying BSIZE
// [from, to): `from` must be the start of a block; scanning stops at the first block starting at or after `to`.
// in_beg is relative to `base` (the file offset the device image starts at).
int scan_bgzf(const uint8_t *f, size_t n, size_t from, size_t to, size_t base, std::vector<BgzfBlock> &blk, std::vector<uint64_t> &coff) {
    size_t off = from;
    unsigned long long out = 0;
    while (off < n && off < to) {
        if (off + 18 > n) return fail(SVTK_ERR_IO, "bam: truncated BGZF header at byte %zu", off);
        const uint8_t *h = f + off;
        if (h[0] != 0x1f || h[1] != 0x8b || h[2] != 8 || !(h[3] & 4)) return fail(SVTK_ERR_IO, "bam: not a BGZF block at byte %zu", off);
        const uint32_t xlen = le16(h + 10);
        if (off + 12 + xlen > n) return fail(SVTK_ERR_IO, "bam: truncated BGZF extra field at byte %zu", off);
        uint32_t bsize = 0;
        for (uint32_t x = 0; x + 4 <= xlen;) {
            const uint8_t *e = h + 12 + x;
            const uint32_t slen = le16(e + 2);
            if (e[0] == 'B' && e[1] == 'C' && slen == 2 && x + 6 <= xlen) bsize = le16(e + 4) + 1;
            x += 4 + slen;
        }
        if (bsize < xlen + 20 || off + bsize > n) return fail(SVTK_ERR_IO, "bam: bad BGZF block size at byte %zu", off);
        BgzfBlock b;
        b.in_beg = off - base + 12 + xlen; b.in_len = bsize - xlen - 20; b.isize = le32(h + bsize - 4); b.out_off = out;
        if (b.isize > 65536u) return fail(SVTK_ERR_IO, "bam: BGZF block at byte %zu claims %u inflated bytes", off, b.isize);
        blk.push_back(b); coff.push_back(off);
        out += b.isize;
        off += bsize;
    }
    return 0;
}

// the BAM header (magic, text, reference dictionary): inflated on the host with the same decoder, block by block, until complete
int read_header(const uint8_t *f, size_t n, std::vector<std::string> &names, std::vector<uint32_t> &lens, unsigned long long *first_rec) {
    std::vector<BgzfBlock> blk;                           // the header's own blocks, scanned lazily from the start of the file
    std::vector<uint64_t> hcoff;
    size_t scanned = 0;
    std::vector<uint32_t> words;                          // the decoder reads aligned words: one block's payload at a time
    std::vector<uint8_t> hd;
    size_t nb = 0;
    std::vector<uint16_t> tabs((1 << BZ_LBITS) + (1 << BZ_DBITS));
    BzTables T;
    T.lit = tabs.data(); T.dist = tabs.data() + (1 << BZ_LBITS);
    auto more = [&](size_t need) -> int {
        while (hd.size() < need) {
            if (nb >= blk.size()) {
                if (scanned >= n) return fail(SVTK_ERR_IO, "bam: header runs past the end of the file");
                const size_t before = blk.size();
                if (int rc = scan_bgzf(f, n, scanned, std::min(n, scanned + (1u << 20)), 0, blk, hcoff)) return rc;
                if (blk.size() == before) return fail(SVTK_ERR_IO, "bam: header runs past the end of the file");
                scanned = (size_t)(blk.back().in_beg + blk.back().in_len + 8);      // base 0: in_beg is a file offset
            }
            const BgzfBlock &b = blk[nb++];
            words.assign((size_t)(b.in_len + BZ_OVERREAD) / 4 + 4, 0u);
            memcpy(words.data(), f + b.in_beg, b.in_len);
            const size_t at = hd.size();
            hd.resize(at + b.isize);
            uint8_t dummy = 0;
            const int rc = bz_inflate(words.data(), 0, b.in_len, b.isize ? hd.data() + at : &dummy, b.isize, T);
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

// record-aligned cut points from the .bai (SAMv1 5.2): linear-index offsets and the chunk starts of every real bin
struct BaiRef { uint64_t beg = ~0ull, end = 0; };          // virtual offsets: first record / past the last record of one reference

int read_bai(const char *path, int *n_ref_out, std::vector<uint64_t> &voffs, std::vector<BaiRef> &refs) {
    Mapped m;
    if (m.open_file(path) != 0) return fail(SVTK_ERR_IO, "bam: cannot open index %s (the reference needs it too: sam_index_load, audit.c:272)", path);
    const uint8_t *f = m.p;
    const size_t n = m.n;
    if (n < 8 || memcmp(f, "BAI\1", 4) != 0) return fail(SVTK_ERR_IO, "bam: %s is not a BAI index", path);
    const int32_t n_ref = (int32_t)le32(f + 4);
    if (n_ref < 0) return fail(SVTK_ERR_IO, "bam: bad index");
    *n_ref_out = n_ref;
    refs.assign((size_t)n_ref, BaiRef());
    size_t p = 8;
    for (int r = 0; r < n_ref; ++r) {
        if (p + 4 > n) return fail(SVTK_ERR_IO, "bam: truncated index");
        const int32_t n_bin = (int32_t)le32(f + p);
        p += 4;
        for (int b = 0; b < n_bin; ++b) {
            if (p + 8 > n) return fail(SVTK_ERR_IO, "bam: truncated index");
            const uint32_t bin = le32(f + p);
            const int32_t n_chunk = (int32_t)le32(f + p + 4);
            p += 8;
            if (n_chunk < 0 || p + 16ull * (size_t)n_chunk > n) return fail(SVTK_ERR_IO, "bam: truncated index");
            if (bin != 37450u)                                      // the pseudo-bin holds file-level metadata, not record offsets
                for (int k = 0; k < n_chunk; ++k) {
                    const uint64_t vb = le64(f + p + 16 * (size_t)k), ve = le64(f + p + 16 * (size_t)k + 8);
                    voffs.push_back(vb);
                    refs[(size_t)r].beg = std::min(refs[(size_t)r].beg, vb); refs[(size_t)r].end = std::max(refs[(size_t)r].end, ve);
                }
            p += 16 * (size_t)n_chunk;
        }
        if (p + 4 > n) return fail(SVTK_ERR_IO, "bam: truncated index");
        const int32_t n_intv = (int32_t)le32(f + p);
        p += 4;
        if (n_intv < 0 || p + 8ull * (size_t)n_intv > n) return fail(SVTK_ERR_IO, "bam: truncated index");
        for (int k = 0; k < n_intv; ++k) { const uint64_t v = le64(f + p + 8 * (size_t)k); if (v) voffs.push_back(v); }
        p += 8 * (size_t)n_intv;
    }
    return 0;
}

// host -> device copy of a (pageable, memory-mapped) file image: the ctx's two pinned staging buffers, each chunk copied into
// staging by several threads at once (one thread faults pages in at ~10 GB/s, the DMA engine takes ~50 GB/s)
int upload_file(Ctx *c, void *dst, const uint8_t *src, size_t bytes) {
    const size_t CH = 32u << 20;
    for (int i = 0; i < 2; ++i) {
        if (int rc = c->stage[i].reserve(CH)) return rc;
        if (!c->stage_ev[i]) CU(cudaEventCreateWithFlags(&c->stage_ev[i], cudaEventDisableTiming));
    }
    const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
    const int T = (int)std::min(8u, std::max(1u, hw / 2));
    size_t done = 0;
    int k = 0;
    while (done < bytes) {
        const size_t n = std::min(CH, bytes - done);
        CU(cudaEventSynchronize(c->stage_ev[k]));
        char *st = (char *)c->stage[k].p;
        const size_t per = ((n + T - 1) / T + 4095) & ~(size_t)4095;
        std::vector<std::thread> th;
        for (int t = 1; t < T; ++t) {
            const size_t a = per * t;
            if (a < n) th.emplace_back([=]() { memcpy(st + a, src + done + a, std::min(per, n - a)); });
        }
        memcpy(st, src + done, std::min(per, n));
        for (auto &x : th) x.join();
        CU(cudaMemcpyAsync((char *)dst + done, st, n, cudaMemcpyHostToDevice, c->stream));
        CU(cudaEventRecord(c->stage_ev[k], c->stream));
        done += n;
        k ^= 1;
    }
    return 0;
}

double ms_since(std::chrono::steady_clock::time_point t0) {
    return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
}

}  // namespace

static int load_bam_impl(svtk_ctx *ctx, const char *bam_path, const char *bai_path, int tid_lo, int tid_hi, svtk_bam_stats *stats) {
    Ctx *c = reinterpret_cast<Ctx *>(ctx);
    if (!c || !bam_path) return fail(SVTK_ERR_ARG, "null argument");
    if (stats) memset(stats, 0, sizeof(*stats));
    if (int rc = set_device(c)) return rc;
    const auto t_all = std::chrono::steady_clock::now();
    c->have_reads = false;
    const bool ranged = tid_lo >= 0;

    // ---- host: map the file; the index first (it says which bytes of the file a reference range needs)
    Mapped bam;
    if (bam.open_file(bam_path) != 0) return fail(SVTK_ERR_IO, "bam: cannot open %s", bam_path);
    if (bam.n == 0) return fail(SVTK_ERR_IO, "bam: %s is empty", bam_path);
    if (bam.n >= (1ull << 34) - 65536) return fail(SVTK_ERR_ARG, "bam: files of 16 GiB and more must be sharded by region (32-bit word index in the decoder)");
    std::vector<uint64_t> voffs;
    std::vector<BaiRef> bai_refs;
    int bai_n_ref = 0;
    {
        std::string bai = bai_path ? std::string(bai_path) : std::string(bam_path) + ".bai";
        if (!bai_path && access(bai.c_str(), R_OK) != 0) {
            std::string alt(bam_path);
            if (alt.size() > 4 && alt.compare(alt.size() - 4, 4, ".bam") == 0) { alt.replace(alt.size() - 4, 4, ".bai"); if (access(alt.c_str(), R_OK) == 0) bai = alt; }
        }
        if (int rc = read_bai(bai.c_str(), &bai_n_ref, voffs, bai_refs)) return rc;
    }
    // byte range [fb, fe) of the file to ship and record range [v_beg, v_end) (virtual offsets) to walk; whole file: everything
    size_t fb = 0, fe = bam.n;
    uint64_t v_beg = 0, v_end = ~0ull;
    bool nothing = false;
    if (ranged) {
        if (tid_hi < tid_lo || tid_hi >= bai_n_ref) return fail(SVTK_ERR_ARG, "bam: reference range [%d, %d] outside the index's %d references", tid_lo, tid_hi, bai_n_ref);
        v_beg = ~0ull; v_end = 0;
        for (int t = tid_lo; t <= tid_hi; ++t) { v_beg = std::min(v_beg, bai_refs[(size_t)t].beg); v_end = std::max(v_end, bai_refs[(size_t)t].end); }
        if (v_beg >= v_end) nothing = true;                        // no reads on these references
        else {
            fb = (size_t)(v_beg >> 16);
            fe = (size_t)(v_end >> 16) + 1;                        // the block holding v_end is the last one scanned (scan stops at the first block at or after fe)
            if (fb >= bam.n) return fail(SVTK_ERR_IO, "bam: index offset beyond the end of the file (stale .bai?)");
        }
    }

    // the table work runs on a helper thread while this thread feeds the upload ring (both only read the mapping)
    std::vector<BgzfBlock> blk;
    std::vector<uint64_t> coff;
    std::vector<unsigned long long> splits;
    unsigned long long total = 0, first_rec = 0, last_rec = 0;
    int n_ref = 0, n_seg = 0, prep_rc = 0;
    std::string prep_msg;
    double prep_ms = 0;
    size_t image_bytes = 0;                                        // bytes of the file the device image holds, from fb
    auto to_linear = [&](uint64_t v, unsigned long long *lin) -> int {
        const uint64_t co = v >> 16, uo = v & 0xffffu;
        auto it = std::lower_bound(coff.begin(), coff.end(), co);
        if (it == coff.end() || *it != co) {
            if (!coff.empty() && co == coff.back() + (blk.back().in_beg + fb - coff.back()) + blk.back().in_len + 8 && uo == 0) { *lin = total; return 0; }   // just past the last block
            return fail(SVTK_ERR_IO, "bam: index offset %llu does not point at a BGZF block (stale .bai?)", (unsigned long long)co);
        }
        const BgzfBlock &b = blk[(size_t)(it - coff.begin())];
        if (uo > b.isize) return fail(SVTK_ERR_IO, "bam: index offset beyond its block (stale .bai?)");
        *lin = b.out_off + uo;
        return 0;
    };
    auto prep = [&]() -> int {
        unsigned long long hdr_end = 0;
        if (int rc = read_header(bam.p, bam.n, c->bam_ref_names, c->bam_ref_lens, &hdr_end)) return rc;
        n_ref = (int)c->bam_ref_names.size();
        if (n_ref != bai_n_ref) return fail(SVTK_ERR_IO, "bam: index has %d references, BAM header %d", bai_n_ref, n_ref);
        if (nothing) { splits.push_back(0); return 0; }
        if (int rc = scan_bgzf(bam.p, bam.n, fb, fe, fb, blk, coff)) return rc;
        if (blk.empty()) return fail(SVTK_ERR_IO, "bam: no BGZF block at index offset %zu (stale .bai?)", fb);
        total = blk.back().out_off + blk.back().isize;
        if (blk.size() >= 0xfffffff0ull) return fail(SVTK_ERR_ARG, "bam: more than 2^32 BGZF blocks: shard the file");
        if (ranged) {
            if (int rc = to_linear(v_beg, &first_rec)) return rc;
            if (int rc = to_linear(v_end, &last_rec)) return rc;
        } else { first_rec = hdr_end; last_rec = total; }
        if (first_rec > last_rec || last_rec > total) return fail(SVTK_ERR_IO, "bam: header / index offsets beyond the file");
        splits.reserve(voffs.size() + 2);
        splits.push_back(first_rec);
        for (uint64_t v : voffs) {
            if (ranged && (v <= v_beg || v >= v_end)) continue;
            unsigned long long lin = 0;
            if (int rc = to_linear(v, &lin)) return rc;
            if (lin > first_rec && lin < last_rec) splits.push_back(lin);
        }
        std::sort(splits.begin(), splits.end());
        splits.erase(std::unique(splits.begin(), splits.end()), splits.end());
        n_seg = first_rec < last_rec ? (int)splits.size() : 0;
        splits.push_back(last_rec);
        return 0;
    };
    std::thread prep_thread([&]() {
        prep_rc = prep();
        if (prep_rc) prep_msg = svtk_last_error();            // the message lives in the helper thread's slot
        prep_ms = ms_since(t_all);
    });
    struct Joiner { std::thread &t; ~Joiner() { if (t.joinable()) t.join(); } } joiner{prep_thread};

    // ---- device: the compressed image (needs nothing from the tables).  A ranged load does not know yet where its last block
    // ends: it ships up to the end of the block that holds v_end, which is at most 64 KiB past fe
    for (int i = 0; i < 4; ++i) if (!c->bam_ev[i]) CU(cudaEventCreate(&c->bam_ev[i]));
    image_bytes = nothing ? 0 : std::min(bam.n, ranged ? fe + 65536 : bam.n) - fb;
    if (int rc = c->bam_comp.reserve(image_bytes + BZ_OVERREAD)) return rc;
    CU(cudaMemsetAsync((char *)c->bam_comp.p + (image_bytes & ~(size_t)3), 0, BZ_OVERREAD - (image_bytes & 3), c->stream));   // the decoder looks a few bytes past the end
    CU(cudaEventRecord(c->bam_ev[0], c->stream));
    if (image_bytes) if (int rc = upload_file(c, c->bam_comp.p, bam.p + fb, image_bytes)) return rc;
    prep_thread.join();
    if (prep_rc) return fail(prep_rc, "%s", prep_msg.c_str());
    const double host_ms = prep_ms;

    // ---- device: tables, inflate
    if (int rc = c->bam_infl.reserve((size_t)total + 64)) return rc;
    if (int rc = c->bam_blocks.reserve(sizeof(BgzfBlock) * std::max<size_t>(blk.size(), 1))) return rc;
    if (int rc = c->bam_splits.reserve(8 * splits.size())) return rc;
    if (int rc = c->bam_seg.reserve(16 * (size_t)std::max(n_seg, 1))) return rc;
    if (int rc = c->bam_ctl.reserve(256)) return rc;
    if (int rc = c->bam_h.reserve(256 + 8 * (size_t)(n_ref + 1) + 16 * (size_t)std::max(n_seg, 1))) return rc;
    CU(cudaMemsetAsync(c->bam_ctl.p, 0, 256, c->stream));
    CU(cudaMemsetAsync((char *)c->bam_infl.p + (total & ~3ull), 0, 64 - (total & 3), c->stream));
    if (int rc = upload(c, c->bam_blocks.p, blk.data(), sizeof(BgzfBlock) * blk.size())) return rc;
    if (int rc = upload(c, c->bam_splits.p, splits.data(), 8 * splits.size())) return rc;
    CU(cudaEventRecord(c->bam_ev[1], c->stream));
    unsigned int *ctl = (unsigned int *)c->bam_ctl.p;
    const uint32_t *infl = (const uint32_t *)c->bam_infl.p;
    if (!blk.empty()) {
        int sms = 148;
        CU(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, c->device));
        // thin warps while the blocks are few (see k_bgzf_inflate): measured best at 12-16 warps per SM (B200: 85 MB / 238 MB / 950 MB
        // files 2.9x / 2.0x / 1.2x faster than with 32 lanes), then widen
        long long want_warps = (long long)sms * 20;
        if (const char *e = getenv("SVTK_BZ_WARPS")) if (*e) want_warps = std::max(1ll, atoll(e));
        int lanes = 1;
        while (lanes < 32 && (long long)((blk.size() + lanes - 1) / lanes) > want_warps) lanes <<= 1;
        const int smem = lanes * BZ_TABLE_WORDS * 4;
        static unsigned long long attr_seen = 0;
        if (first_use_on_device(attr_seen)) CU(cudaFuncSetAttribute(k_bgzf_inflate, cudaFuncAttributeMaxDynamicSharedMemorySize, 32 * BZ_TABLE_WORDS * 4));
        int per_sm = 1;
        CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_bgzf_inflate, 32, smem));
        const unsigned grid = (unsigned)std::min<size_t>((size_t)sms * std::max(per_sm, 1), (blk.size() + lanes - 1) / lanes);
        k_bgzf_inflate<<<grid, 32, smem, c->stream>>>((const uint32_t *)c->bam_comp.p, (const BgzfBlock *)c->bam_blocks.p, (unsigned)blk.size(),
                                                     (uint8_t *)c->bam_infl.p, ctl, lanes);
        if (stats) stats->lanes_per_warp = (uint32_t)lanes;
        CU(cudaGetLastError());
        c->launches += 1;
    }
    CU(cudaEventRecord(c->bam_ev[2], c->stream));
    if (!blk.empty()) {
        int sms = 148;
        CU(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, c->device));
        const unsigned grid = (unsigned)std::min<size_t>((size_t)sms * 8, (blk.size() + 7) / 8);
        k_bgzf_crc<<<grid, 256, 0, c->stream>>>((const uint32_t *)c->bam_comp.p, (const BgzfBlock *)c->bam_blocks.p, (unsigned)blk.size(),
                                              (const uint8_t *)c->bam_infl.p, ctl);
        CU(cudaGetLastError());
        c->launches += 1;
    }

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
    if (n_reads >= (1ll << 31) - 1024) return fail(SVTK_ERR_ARG, "bam: more than 2^31 reads per device: shard the file");
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
    CU(cudaEventRecord(c->bam_ev[3], c->stream));
    std::vector<int64_t> tid_off((size_t)n_ref + 1, 0);
    CU(cudaMemcpyAsync(hctl, ctl, 64, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaMemcpyAsync(hctl + 16, c->tmp_b.p, 8 * (size_t)(n_ref + 1), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    if (hctl[BAM_CTL_ORDER]) return fail(SVTK_ERR_IO, "bam: records are not in coordinate order (or a refID is outside the header's dictionary)");
    memcpy(tid_off.data(), hctl + 16, 8 * (size_t)(n_ref + 1));
    c->bam_infl_bytes = total; c->bam_n_rec = n_reads;
    c->h2d_bytes += image_bytes + sizeof(BgzfBlock) * blk.size() + 8 * splits.size();

    svtk_reads d{};
    d.n_tid = n_ref; d.n_reads = n_reads; d.n_words = n_words; d.tid_off = tid_off.data();
    d.pos = (const uint32_t *)c->r_pos.p; d.end = (const uint32_t *)c->r_end.p; d.n_cigar = (const uint32_t *)c->r_ncig.p;
    d.cigar_off = (const uint64_t *)c->r_off.p; d.cigar = (const uint32_t *)c->r_cigar.p;
    if (int rc = adopt_device_reads(c, &d)) return rc;
    c->reads_from_bam = true;
    if (stats) {
        float up = 0, inf = 0, par = 0;
        cudaEventElapsedTime(&up, c->bam_ev[0], c->bam_ev[1]); cudaEventElapsedTime(&inf, c->bam_ev[1], c->bam_ev[2]); cudaEventElapsedTime(&par, c->bam_ev[2], c->bam_ev[3]);
        stats->file_bytes = bam.n; stats->image_bytes = image_bytes; stats->n_blocks = blk.size(); stats->inflated_bytes = total; stats->n_reads = (uint64_t)n_reads;
        stats->n_words = (uint64_t)n_words; stats->n_ref = n_ref; stats->n_pieces = (uint32_t)n_seg; stats->host_scan_ms = (float)host_ms;
        stats->upload_ms = up; stats->inflate_ms = inf; stats->parse_ms = par; stats->total_ms = (float)ms_since(t_all);
    }
    return 0;
}

extern "C" int svtk_load_bam(svtk_ctx *ctx, const char *bam_path, const char *bai_path, svtk_bam_stats *stats) {
    return load_bam_impl(ctx, bam_path, bai_path, -1, -1, stats);
}

extern "C" int svtk_load_bam_refs(svtk_ctx *ctx, const char *bam_path, const char *bai_path, int32_t tid_lo, int32_t tid_hi, svtk_bam_stats *stats) {
    if (tid_lo < 0) return fail(SVTK_ERR_ARG, "bad reference range");
    return load_bam_impl(ctx, bam_path, bai_path, tid_lo, tid_hi, stats);
}

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
