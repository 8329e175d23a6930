// svtk_bgzf_core.cuh -- RFC 1951 (DEFLATE) decoder for one BGZF block, written so that ONE GPU LANE decodes ONE block
// (32 independent blocks per warp) and so that the same source compiles as plain C++ for the CPU unit tests
// (tests/hostsim/bgzf_hostsim.cpp).
//
// Where it sits: the reference reaches every read through HTSlib, which inflates each 64 KiB BGZF block with zlib on the
// host every time an iterator touches it (bgzf.c:746 bgzf_uncompress -> inflate, called from bgzf_read_block bgzf.c:988);
// >= 99 % of `svtrek audt` wall time is spent there.  BGZF blocks are independent gzip members
// (SAMv1 4.1), so svtk_load_bam (svtk_bam.cu) ships the raw file to HBM and inflates all blocks at once.
//
// Decoder shape (per lane, state in registers, tables in shared memory):
//   - 64-bit LSB-first bit buffer refilled 32 bits at a time from the (word aligned) compressed buffer;
//   - canonical Huffman codes: a primary lookup table (9 bits literal/length, 8 bits distance; entry = symbol << 4 | code
//     length) in shared memory, filled per DEFLATE block, and for the rare longer codes the count/first-code walk over the
//     code lengths (those small arrays live in thread-local memory);
//   - length / distance bases computed arithmetically (no constant-memory tables: lanes index them divergently);
//   - every output byte is bounds-checked against the block's ISIZE, every match distance against the bytes produced so far
//     (a BGZF block has no preset dictionary), so corrupt input yields an error code, never a stray store.
#pragma once
#include <stdint.h>
#include <stddef.h>

#if defined(__CUDACC__)
#define BZ_HD __host__ __device__ __forceinline__
#else
#define BZ_HD inline
#endif

namespace svtk {

#ifndef SVTK_BZ_LBITS
#define SVTK_BZ_LBITS 8
#endif
#ifndef SVTK_BZ_DBITS
#define SVTK_BZ_DBITS 7
#endif
constexpr int BZ_LBITS = SVTK_BZ_LBITS, BZ_DBITS = SVTK_BZ_DBITS;     // primary table bits (>= 7: the code-length code borrows the distance table)
enum { BZ_OK = 0, BZ_ERR_BTYPE = 1, BZ_ERR_STORED = 2, BZ_ERR_LENS = 3, BZ_ERR_CODE = 4, BZ_ERR_DIST = 5, BZ_ERR_OUT = 6,
       BZ_ERR_IN = 7, BZ_ERR_SIZE = 8, BZ_ERR_HEADER = 9 };

struct BzTables {                         // one per lane
    uint16_t *lit;                        // primary literal/length table [1 << BZ_LBITS]: sym << 4 | len, 0 = code longer than BZ_LBITS
    uint16_t *dist;                       // primary distance table [1 << BZ_DBITS]   (both in shared memory on the GPU)
    uint16_t lsym[288], dsym[32];         // symbols in canonical order, codes per length: table builds and the long-code walk only
    uint16_t lcnt[16], dcnt[16];          //   (thread-local memory on the GPU)
    uint32_t lresume, dresume;            // long-code walk resumed after the primary bits: first code << 16 | symbol index
};
constexpr int BZ_TABLE_WORDS = (((1 << BZ_LBITS) + (1 << BZ_DBITS)) / 2) | 1;   // 32-bit words per lane; odd stride: 32 lanes, 32 banks

struct BzBits {
    const uint32_t *w;                    // compressed buffer, 4-byte aligned, readable 8 bytes past the last block
    unsigned long long end;               // byte offset of the end of this block's DEFLATE payload
    uint32_t wi;                          // index of the word held in nxt: everything before it is in buf or consumed (files < 16 GiB)
    unsigned long long buf;
    int cnt;
    uint32_t nxt;                         // the word at ip, loaded one refill ahead
};

BZ_HD uint32_t bz_load32(const uint32_t *w, unsigned long long off) {
    const unsigned long long i = off >> 2;
    const uint32_t sh = (uint32_t)(off & 3) * 8u, lo = w[i], hi = w[i + 1];
#if defined(__CUDA_ARCH__)
    return __funnelshift_r(lo, hi, sh);
#else
    return sh ? (lo >> sh) | (hi << (32u - sh)) : lo;
#endif
}
BZ_HD uint32_t bz_byte(const uint32_t *w, unsigned long long off) { return (w[off >> 2] >> ((uint32_t)(off & 3) * 8u)) & 255u; }

// ip is kept 4-byte aligned so that a refill is ONE aligned load (lanes read different lines: every load instruction costs
// the warp 32 cache accesses, so halving the loads matters more than the shift it costs)
BZ_HD void bz_prime(BzBits &b, unsigned long long pos) {      // start reading bits at byte offset pos
    const uint32_t drop = (uint32_t)(pos & 3) * 8u;
    b.wi = (uint32_t)(pos >> 2) + 1u;
    b.buf = (unsigned long long)(b.w[b.wi - 1u] >> drop); b.cnt = 32 - (int)drop;
    b.nxt = b.w[b.wi];
}
BZ_HD unsigned long long bz_ip(const BzBits &b) { return (unsigned long long)b.wi * 4ull; }   // byte offset of the first byte not yet in buf
BZ_HD void bz_refill(BzBits &b) {         // afterwards 33 <= cnt <= 64; the load issued here is consumed by the NEXT refill
    if (b.cnt <= 32) { b.buf |= (unsigned long long)b.nxt << b.cnt; b.cnt += 32; b.nxt = b.w[++b.wi]; }
}
BZ_HD uint32_t bz_take(BzBits &b, int n) {   // n <= 16, cnt >= n
    const uint32_t v = (uint32_t)b.buf & ((1u << n) - 1u);
    b.buf >>= n; b.cnt -= n;
    return v;
}

BZ_HD uint32_t bz_rev(uint32_t code, int len) {
#if defined(__CUDA_ARCH__)
    return __brev(code) >> (32 - len);
#else
    uint32_t r = 0;
    for (int i = 0; i < len; ++i) r |= ((code >> i) & 1u) << (len - 1 - i);
    return r;
#endif
}

// canonical code from lens[0..n): counts, symbols in code order, primary table of `bits` bits.  Over-subscribed -> error.
BZ_HD int bz_build(const uint8_t *lens, int n, uint16_t *cnt, uint16_t *sym, uint16_t *tab, int bits, uint32_t *resume) {
    for (int l = 0; l < 16; ++l) cnt[l] = 0;
    for (int s = 0; s < n; ++s) cnt[lens[s]]++;
    int left = 1;
    for (int l = 1; l < 16; ++l) { left = (left << 1) - (int)cnt[l]; if (left < 0) return BZ_ERR_LENS; }
    uint16_t offs[16];
    offs[0] = 0; offs[1] = 0;
    for (int l = 1; l < 15; ++l) offs[l + 1] = (uint16_t)(offs[l] + cnt[l]);
    for (int s = 0; s < n; ++s) if (lens[s]) sym[offs[lens[s]]++] = (uint16_t)s;
    for (int i = 0; i < (1 << bits); ++i) tab[i] = 0;
    uint32_t code = 0;
    int idx = 0;
    for (int l = 1; l <= bits; ++l) {
        for (int k = 0; k < (int)cnt[l]; ++k, ++idx, ++code) {
            const uint16_t e = (uint16_t)(sym[idx] << 4 | l);
            for (uint32_t r = bz_rev(code, l); r < (1u << bits); r += 1u << l) tab[r] = e;
        }
        code <<= 1;
    }
    *resume = code << 16 | (uint32_t)idx;                   // state of the long-code walk after `bits` lengths
    return BZ_OK;
}

// one symbol: primary table, else the bit-by-bit walk (codes longer than `bits`).  cnt >= 15 bits available.  -1 = invalid code
BZ_HD int bz_symbol(BzBits &b, const uint16_t *tab, int bits, const uint16_t *cnt, const uint16_t *sym, uint32_t resume) {
    const uint32_t e = tab[(uint32_t)b.buf & ((1u << bits) - 1u)];
    if (e) { const int l = (int)(e & 15u); b.buf >>= l; b.cnt -= l; return (int)(e >> 4); }
    // codes longer than the primary table: the canonical walk (first code / count per length), resumed at bits + 1
    int code = (int)(bz_rev((uint32_t)b.buf & ((1u << bits) - 1u), bits) << 1), first = (int)(resume >> 16), index = (int)(resume & 0xffffu);
    unsigned long long v = b.buf >> bits;
    for (int l = bits + 1; l < 16; ++l) {
        code |= (int)(v & 1ull); v >>= 1;
        const int c = (int)cnt[l];
        if (code - c < first) { b.buf >>= l; b.cnt -= l; return (int)sym[index + (code - first)]; }
        index += c; first += c; first <<= 1; code <<= 1;
    }
    return -1;
}

BZ_HD uint32_t bz_clen_order(int i) {    // RFC 1951 3.2.7: 16,17,18,0,8,7,9,6,10,5,11,4 | 12,3,13,2,14,1,15 (5 bits each)
    const unsigned long long a = 16ull | 17ull << 5 | 18ull << 10 | 0ull << 15 | 8ull << 20 | 7ull << 25 | 9ull << 30 | 6ull << 35 |
                                 10ull << 40 | 5ull << 45 | 11ull << 50 | 4ull << 55;
    const unsigned long long c = 12ull | 3ull << 5 | 13ull << 10 | 2ull << 15 | 14ull << 20 | 1ull << 25 | 15ull << 30;
    return (uint32_t)((i < 12 ? a >> (5 * i) : c >> (5 * (i - 12))) & 31ull);
}

// ---- resumable per-lane decoder.  On the GPU the 32 lanes of a warp run bz_header / bz_symbols from ONE warp-level loop
// (svtk_bam.cu k_bgzf_inflate) so that lanes working on different blocks stay converged on the symbol step; a lane that
// finished its block must not drift into a private copy of the loop.  On the host, bz_inflate below drives one lane.
enum { BZ_S_IDLE = 0, BZ_S_HEADER = 1, BZ_S_SYMBOLS = 2 };

struct BzLane {
    BzBits b;
    uint8_t *out;
    uint32_t op, out_len;                 // bytes produced so far / expected (ISIZE)
    unsigned long long *tok;              // match tokens of this block (bz_step queues, bz_resolve_* copies): pos | len << 16 | (dist - 1) << 32
    uint32_t n_tok;
    int state, last, err;
};

BZ_HD void bz_begin(BzLane &z, const uint32_t *words, unsigned long long in_beg, unsigned long long in_end, uint8_t *out, uint32_t out_len,
                     unsigned long long *tok) {
    z.b.w = words; z.b.end = in_end;
    bz_prime(z.b, in_beg);
    z.out = out; z.op = 0; z.out_len = out_len; z.tok = tok; z.n_tok = 0; z.state = BZ_S_HEADER; z.last = 0; z.err = BZ_OK;
}

BZ_HD void bz_fail(BzLane &z, int e) { z.err = e; z.state = BZ_S_IDLE; }

BZ_HD void bz_finish(BzLane &z) {        // after the final block
    if (bz_ip(z.b) - (unsigned long long)(z.b.cnt >> 3) > z.b.end) z.err = BZ_ERR_IN;
    else if (z.op != z.out_len) z.err = BZ_ERR_SIZE;
    z.state = BZ_S_IDLE;
}

// one DEFLATE block header: stored blocks are copied here; for Huffman blocks the lane's tables are (re)built
BZ_HD void bz_header(BzLane &z, BzTables &T) {
    BzBits &b = z.b;
    if (bz_ip(b) > b.end + 8) return bz_fail(z, BZ_ERR_IN);
    bz_refill(b);
    z.last = (int)bz_take(b, 1);
    const uint32_t type = bz_take(b, 2);
    if (type == 3) return bz_fail(z, BZ_ERR_BTYPE);
    if (type == 0) {                                                       // stored: byte align, LEN, ~LEN, bytes
        bz_take(b, b.cnt & 7);
        unsigned long long ip = bz_ip(b) - (unsigned long long)(b.cnt >> 3);   // hand the whole bytes back
        if (ip + 4 > b.end) return bz_fail(z, BZ_ERR_STORED);
        const uint32_t len = bz_byte(b.w, ip) | bz_byte(b.w, ip + 1) << 8;
        const uint32_t nlen = bz_byte(b.w, ip + 2) | bz_byte(b.w, ip + 3) << 8;
        ip += 4;
        if ((len ^ nlen) != 0xffffu || ip + len > b.end) return bz_fail(z, BZ_ERR_STORED);
        if (z.op + len > z.out_len) return bz_fail(z, BZ_ERR_OUT);
        for (uint32_t k = 0; k < len; ++k) z.out[z.op + k] = (uint8_t)bz_byte(b.w, ip + k);
        z.op += len;
        bz_prime(b, ip + len);
        if (z.last) bz_finish(z);
        return;
    }
    uint8_t lens[320];
    int hlit, hdist;
    if (type == 1) {                                                       // fixed code, RFC 1951 3.2.6
        hlit = 288; hdist = 30;
        for (int s = 0; s < 144; ++s) lens[s] = 8;
        for (int s = 144; s < 256; ++s) lens[s] = 9;
        for (int s = 256; s < 280; ++s) lens[s] = 7;
        for (int s = 280; s < 288; ++s) lens[s] = 8;
        for (int s = 0; s < 30; ++s) lens[288 + s] = 5;
    } else {
        hlit = (int)bz_take(b, 5) + 257; hdist = (int)bz_take(b, 5) + 1;
        const int hclen = (int)bz_take(b, 4) + 4;
        if (hlit > 286 || hdist > 30) return bz_fail(z, BZ_ERR_HEADER);
        for (int i = 0; i < 19; ++i) lens[i] = 0;
        for (int i = 0; i < hclen; ++i) { bz_refill(b); lens[bz_clen_order(i)] = (uint8_t)bz_take(b, 3); }
        // the code-length code borrows the distance table's storage (its 7-bit primary table covers every code)
        if (bz_build(lens, 19, T.dcnt, T.dsym, T.dist, 7, &T.dresume)) return bz_fail(z, BZ_ERR_LENS);
        int n = 0, prev = 0;
        while (n < hlit + hdist) {
            bz_refill(b);
            const int s = bz_symbol(b, T.dist, 7, T.dcnt, T.dsym, T.dresume);
            if (s < 0) return bz_fail(z, BZ_ERR_CODE);
            if (s < 16) { lens[n++] = (uint8_t)s; prev = s; continue; }
            int rep, val = 0;
            if (s == 16) { if (n == 0) return bz_fail(z, BZ_ERR_LENS); val = prev; rep = 3 + (int)bz_take(b, 2); }
            else if (s == 17) { rep = 3 + (int)bz_take(b, 3); prev = 0; }
            else { rep = 11 + (int)bz_take(b, 7); prev = 0; }
            if (n + rep > hlit + hdist) return bz_fail(z, BZ_ERR_LENS);
            while (rep--) lens[n++] = (uint8_t)val;
        }
        if (lens[256] == 0) return bz_fail(z, BZ_ERR_LENS);                // no end-of-block code
        if (bz_ip(b) > b.end + 8) return bz_fail(z, BZ_ERR_IN);
    }
    if (bz_build(lens, hlit, T.lcnt, T.lsym, T.lit, BZ_LBITS, &T.lresume)) return bz_fail(z, BZ_ERR_LENS);
    if (bz_build(lens + hlit, hdist, T.dcnt, T.dsym, T.dist, BZ_DBITS, &T.dresume)) return bz_fail(z, BZ_ERR_LENS);
    z.state = BZ_S_SYMBOLS;
}

// ---- length / distance symbol -> {base, extra bits}.  On the GPU two small tables in shared memory (lanes of a warp index them
// divergently: 29 / 30 words, one bank each, so no conflicts); arithmetic on the host.  RFC 1951 3.2.5.
BZ_HD uint32_t bz_len_entry(int k) {      // k = symbol - 257 in 0..28 -> base | extra bits << 16
    if (k < 8) return (uint32_t)(k + 3);
    if (k == 28) return 258u;
    const int eb = (k - 4) >> 2;
    return (3u + ((4u + (uint32_t)(k & 3)) << eb)) | (uint32_t)eb << 16;
}
BZ_HD uint32_t bz_dist_entry(int d) {     // d in 0..29 -> base | extra bits << 16
    if (d < 4) return (uint32_t)d + 1u;
    const int eb = (d >> 1) - 1;
    return (1u + ((2u + (uint32_t)(d & 1)) << eb)) | (uint32_t)eb << 16;
}
constexpr int BZ_LUT_WORDS = 64;          // [0..28] lengths, [32..61] distances
struct BzLut { const uint32_t *w; };      // nullptr: compute (host)
BZ_HD uint32_t bz_len_lookup(const BzLut &L, int k) { return L.w ? L.w[k] : bz_len_entry(k); }
BZ_HD uint32_t bz_dist_lookup(const BzLut &L, int d) { return L.w ? L.w[32 + d] : bz_dist_entry(d); }

// ---- LZ77 in two phases.  A lane that copied its matches as it met them waited a DRAM round trip per match (the window of
// every block in flight does not fit L2: 27 % of the kernel's stall samples sat on the first use of the source bytes, B200,
// profiles/r02b) and dragged its warp through copy loops of lane-dependent length.  So the decoding lane only QUEUES a match --
// destination offset, length, distance: one 8-byte token -- and skips the bytes; literals are stored directly.  A second pass
// (k_bgzf_resolve: one WARP per block, 32 matches in flight at a time, full occupancy) fills the holes in stream order.
// A block of n inflated bytes holds at most n / 3 matches.
BZ_HD unsigned long long bz_token(uint32_t pos, uint32_t len, uint32_t dist) {
    return (unsigned long long)(pos | len << 16) | (unsigned long long)(dist - 1u) << 32;   // two 32-bit halves: no shift across the words
}
BZ_HD uint32_t bz_tok_pos(unsigned long long t) { return (uint32_t)t & 0xffffu; }
BZ_HD uint32_t bz_tok_len(unsigned long long t) { return ((uint32_t)t >> 16) & 511u; }
BZ_HD uint32_t bz_tok_dist(unsigned long long t) { return (uint32_t)(t >> 32) + 1u; }
BZ_HD unsigned long long bz_tok_capacity(unsigned long long inflated_bytes, unsigned long long n_blocks) { return inflated_bytes / 3ull + n_blocks; }
BZ_HD unsigned long long bz_tok_offset(unsigned long long out_off, unsigned long long block_index) { return out_off / 3ull + block_index; }

// up to BZ_CHUNK bytes of one match: o[0 .. n) = o[-dist ..], n <= BZ_CHUNK.  dist >= BZ_CHUNK: the source lies wholly behind the
// chunk.  dist < BZ_CHUNK (a run): the period is replicated from a register; *dist is then raised to the smallest multiple of
// the period >= BZ_CHUNK (same bytes: the data is periodic), so the match's later chunks are plain copies again.
constexpr uint32_t BZ_CHUNK = 8;
BZ_HD void bz_copy_chunk(uint8_t *o, uint32_t *dist_io, uint32_t n) {
    const uint32_t dist = *dist_io;
    unsigned long long v;                 // the chunk's bytes, first byte in the low bits
#if defined(__CUDA_ARCH__)
    {   // three aligned words cover the 8 source bytes wherever they start (bytes past the source are ignored; the buffers are
        // readable 16 bytes past their end)
        const uintptr_t a = (uintptr_t)(o - dist);
        const volatile uint32_t *aw = (const volatile uint32_t *)(a & ~(uintptr_t)3);
        const uint32_t sh = (uint32_t)(a & 3) * 8u, w0 = aw[0], w1 = aw[1], w2 = aw[2];
        v = (unsigned long long)__funnelshift_r(w0, w1, sh) | (unsigned long long)__funnelshift_r(w1, w2, sh) << 32;
    }
#else
    v = 0;
    for (uint32_t i = 0; i < BZ_CHUNK && i < dist; ++i) v |= (unsigned long long)o[(long long)i - (long long)dist] << (8u * i);
#endif
    if (dist < BZ_CHUNK) {                // run: bytes [0, dist) of v are the period; replicate it over the chunk
        v &= (1ull << (8u * dist)) - 1ull;
        uint32_t have = dist;
        while (have < BZ_CHUNK) { v |= v << (8u * have); have <<= 1; }      // doubling keeps the phase: have stays a multiple of dist
        *dist_io = (uint32_t)(0x0E0C0A0809080800ull >> (8u * dist)) & 255u;  // 1..7 -> 8 8 9 8 10 12 14
    }
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (uint32_t i = 0; i < BZ_CHUNK; ++i) if (i < n) o[i] = (uint8_t)(v >> (8u * i));
}

// the whole second phase for one block, by one thread (host driver, CPU tests)
BZ_HD void bz_resolve_serial(uint8_t *out, const unsigned long long *tok, uint32_t n_tok) {
    for (uint32_t k = 0; k < n_tok; ++k) {
        const unsigned long long t = tok[k];
        uint32_t dist = bz_tok_dist(t), left = bz_tok_len(t);
        uint8_t *o = out + bz_tok_pos(t);
        while (left) { const uint32_t n = left < BZ_CHUNK ? left : BZ_CHUNK; bz_copy_chunk(o, &dist, n); o += n; left -= n; }
    }
}

// A corrupt stream can keep decoding "symbols" out of whatever follows its payload; the bound is checked once per burst of
// BZ_BURST steps (a step consumes at most 2 x 48 bits), so the compressed buffer must be readable BZ_OVERREAD bytes past its end.
constexpr int BZ_BURST = 64;
constexpr int BZ_OVERREAD = 16 * BZ_BURST + 64;
BZ_HD void bz_check_input(BzLane &z) {
    if (z.state != BZ_S_IDLE && bz_ip(z.b) > z.b.end + 8) bz_fail(z, BZ_ERR_IN);
}

// One round of a lane in state BZ_S_SYMBOLS: up to two literal/length symbols (the second only after a literal: a refill leaves
// >= 32 bits, a literal takes <= 15); literals are stored, a match is range-checked and queued.
BZ_HD void bz_step(BzLane &z, const BzTables &T, const BzLut &L) {
    BzBits &b = z.b;
    bz_refill(b);
    int s = bz_symbol(b, T.lit, BZ_LBITS, T.lcnt, T.lsym, T.lresume);
    if (s < 256 && s >= 0) {
        if (z.op >= z.out_len) return bz_fail(z, BZ_ERR_OUT);
        z.out[z.op++] = (uint8_t)s;
        s = bz_symbol(b, T.lit, BZ_LBITS, T.lcnt, T.lsym, T.lresume);
        if (s < 256 && s >= 0) {
            if (z.op >= z.out_len) return bz_fail(z, BZ_ERR_OUT);
            z.out[z.op++] = (uint8_t)s;
            return;
        }
    }
    if (s < 0) return bz_fail(z, BZ_ERR_CODE);
    if (s == 256) {
        if (z.last) bz_finish(z); else z.state = BZ_S_HEADER;
        return;
    }
    if (s > 285) return bz_fail(z, BZ_ERR_CODE);
    bz_refill(b);
    const uint32_t le = bz_len_lookup(L, s - 257);
    const uint32_t len = (le & 0xffffu) + bz_take(b, (int)(le >> 16));
    const int d = bz_symbol(b, T.dist, BZ_DBITS, T.dcnt, T.dsym, T.dresume);
    if (d < 0 || d > 29) return bz_fail(z, BZ_ERR_CODE);
    const uint32_t de = bz_dist_lookup(L, d);             // no refill: 33 bits cover length extra (5) + distance code (15) + distance extra (13)
    const uint32_t dist = (de & 0xffffu) + bz_take(b, (int)(de >> 16));
    if (dist > z.op) return bz_fail(z, BZ_ERR_DIST);
    if (z.op + len > z.out_len) return bz_fail(z, BZ_ERR_OUT);
    z.tok[z.n_tok++] = bz_token(z.op, len, dist);            // n_tok <= out_len / 3: every match is >= 3 bytes and fits the block
    z.op += len;
}

// Inflate one raw DEFLATE stream (the payload of one BGZF block): words/in_beg/in_end locate it inside the compressed
// buffer, out/out_len is the block's slice of the inflated buffer (out_len == the block's ISIZE footer).  Returns BZ_OK
// when the final block ended with exactly out_len bytes written.  (Single-lane driver; the kernel has its own.)
BZ_HD int bz_inflate(const uint32_t *words, unsigned long long in_beg, unsigned long long in_end, uint8_t *out, uint32_t out_len,
                     BzTables &T, unsigned long long *tok) {          // tok: room for out_len / 3 + 1 tokens
    BzLane z;
    bz_begin(z, words, in_beg, in_end, out, out_len, tok);
    const BzLut L{nullptr};
    while (z.state != BZ_S_IDLE) {
        if (z.state == BZ_S_HEADER) bz_header(z, T);
        else bz_step(z, T, L);
        bz_check_input(z);
    }
    if (z.err == BZ_OK) bz_resolve_serial(out, tok, z.n_tok);
    return z.err;
}

}  // namespace svtk

// ---- CRC-32 of a block's inflated bytes (gzip footer, RFC 1952; HTSlib verifies it for every block it inflates,
// bgzf.c:776-780).  The bytes are cut into up to 32 slices, one per lane; every lane runs the table-driven byte loop on its
// slice, and the slice CRCs are merged with the identity zlib's crc32_combine rests on: appending n bytes to a message
// multiplies its CRC by x^(8n) modulo the CRC polynomial, so crc(A || B) = crc(A) * x^(8|B|) xor crc(B) (zlib 1.2.12+
// crc32.c: multmodp / x2nmodp; restated here, reflected polynomial 0xEDB88320).
namespace svtk {

constexpr uint32_t BZ_CRC_POLY = 0xedb88320u;

BZ_HD uint32_t bz_crc_table_entry(uint32_t i) {             // the classic 256-entry table, one entry
    uint32_t c = i;
    for (int k = 0; k < 8; ++k) c = (c & 1u) ? (c >> 1) ^ BZ_CRC_POLY : c >> 1;
    return c;
}

BZ_HD uint32_t bz_crc_bytes(const uint32_t *tab, const uint8_t *p, uint32_t n) {   // standard CRC-32 of n bytes (init / final xor included)
    uint32_t c = 0xffffffffu;
    for (uint32_t i = 0; i < n; ++i) c = tab[(c ^ p[i]) & 255u] ^ (c >> 8);
    return c ^ 0xffffffffu;
}

BZ_HD uint32_t bz_crc_multmodp(uint32_t a, uint32_t b) {    // a(x) * b(x) mod p(x), bit 31 = x^0 (reflected)
    uint32_t m = 1u << 31, p = 0;
    for (;;) {
        if (a & m) { p ^= b; if ((a & (m - 1u)) == 0u) break; }
        m >>= 1;
        b = (b & 1u) ? (b >> 1) ^ BZ_CRC_POLY : b >> 1;
    }
    return p;
}

BZ_HD uint32_t bz_crc_x8n(uint32_t n) {                     // x^(8 n) mod p(x): square-and-multiply over the bits of n
    uint32_t sq = bz_crc_multmodp(1u << 30, 1u << 30);      // x^2
    sq = bz_crc_multmodp(sq, sq); sq = bz_crc_multmodp(sq, sq);   // x^8
    uint32_t p = 1u << 31;                                  // x^0
    while (n) {
        if (n & 1u) p = bz_crc_multmodp(sq, p);
        n >>= 1;
        if (n) sq = bz_crc_multmodp(sq, sq);
    }
    return p;
}

// CRC of `n` bytes from the CRCs of its slices: slice s covers bytes [s * per, min(n, (s + 1) * per)); this is lane s's share
// (XOR the shares of all slices to get the CRC of the whole)
BZ_HD uint32_t bz_crc_share(uint32_t slice_crc, uint32_t s, uint32_t per, uint32_t n) {
    const unsigned long long endb = (unsigned long long)(s + 1u) * per;
    const uint32_t after = endb >= n ? 0u : n - (uint32_t)endb;
    return after ? bz_crc_multmodp(bz_crc_x8n(after), slice_crc) : slice_crc;
}

}  // namespace svtk
