/* bgzfpack.c -- INPUT-GENERATOR helper (not product code, not the oracle): compresses the BGZF blocks of a synthetic BAM in
 * parallel with zlib, for svtrek_b200/simreads.py:write_bam_fast (bench.py writes a ~2 GB config-2 BAM with it).
 * Block layout per SAMv1 4.1: gzip member with the BC extra subfield, raw DEFLATE payload, CRC32, ISIZE. */
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <zlib.h>

typedef struct {
    const uint8_t *data; const int64_t *off; int n_blocks, level;
    uint8_t *out; int64_t stride; int64_t *out_size;
    int next; pthread_mutex_t mu; int err;
} job_t;

static int pack_one(const uint8_t *src, int64_t n, int level, uint8_t *dst, int64_t cap) {
    static const uint8_t head[16] = {0x1f, 0x8b, 8, 4, 0, 0, 0, 0, 0, 0xff, 6, 0, 'B', 'C', 2, 0};
    z_stream z;
    memset(&z, 0, sizeof z);
    if (deflateInit2(&z, level, Z_DEFLATED, -15, 8, Z_DEFAULT_STRATEGY) != Z_OK) return -1;
    z.next_in = (Bytef *)src; z.avail_in = (uInt)n;
    z.next_out = dst + 18; z.avail_out = (uInt)(cap - 18 - 8);
    if (deflate(&z, Z_FINISH) != Z_STREAM_END) { deflateEnd(&z); return -1; }
    const int64_t clen = (int64_t)z.total_out;
    deflateEnd(&z);
    const int64_t bsize = clen + 26;                  /* whole block */
    if (bsize > 65536) return -1;
    memcpy(dst, head, 16);
    dst[16] = (uint8_t)((bsize - 1) & 255); dst[17] = (uint8_t)((bsize - 1) >> 8);
    const uint32_t crc = (uint32_t)crc32(crc32(0L, Z_NULL, 0), src, (uInt)n), isz = (uint32_t)n;
    uint8_t *f = dst + 18 + clen;
    for (int i = 0; i < 4; ++i) { f[i] = (uint8_t)(crc >> (8 * i)); f[4 + i] = (uint8_t)(isz >> (8 * i)); }
    return (int)bsize;
}

static void *worker(void *p) {
    job_t *j = (job_t *)p;
    for (;;) {
        pthread_mutex_lock(&j->mu);
        const int k = j->next++;
        pthread_mutex_unlock(&j->mu);
        if (k >= j->n_blocks) break;
        const int r = pack_one(j->data + j->off[k], j->off[k + 1] - j->off[k], j->level, j->out + (int64_t)k * j->stride, j->stride);
        if (r < 0) { j->err = 1; j->out_size[k] = 0; } else j->out_size[k] = r;
    }
    return NULL;
}

/* blocks k = 0..n_blocks-1 are data[off[k] .. off[k+1]) (each <= 0xff00 bytes); block k is written at out + k * stride
 * (stride >= 65536 + 64) and its size stored in out_size[k].  Returns 0 on success. */
int bgzfpack(const uint8_t *data, const int64_t *off, int n_blocks, int level, int threads, uint8_t *out, int64_t stride, int64_t *out_size) {
    job_t j = {data, off, n_blocks, level, out, stride, out_size, 0, PTHREAD_MUTEX_INITIALIZER, 0};
    if (threads < 1) threads = 1;
    if (threads > 64) threads = 64;
    pthread_t th[64];
    int made = 0;
    for (int t = 0; t < threads - 1; ++t) if (pthread_create(&th[made], NULL, worker, &j) == 0) ++made;
    worker(&j);
    for (int t = 0; t < made; ++t) pthread_join(th[t], NULL);
    return j.err;
}

/* ---- BAM record bytes of a run of reads (SAMv1 4.2): block_size, core, fixed-width name "r%09d", CIGAR, random SEQ / QUAL.
 * Per-read generator state derives from (seed, read index), so the bytes do not depend on the thread count. */
typedef struct {
    int64_t n, first_index; const int32_t *tid; const uint32_t *pos; const uint16_t *bin; const uint32_t *ncig; const int64_t *cig_off;
    const uint32_t *cigar; const int64_t *qlen; const int64_t *rec_off; uint64_t seed; uint8_t *out;
    int64_t next; pthread_mutex_t mu;
} rec_job_t;

static inline uint64_t sm64(uint64_t *s) { uint64_t z = (*s += 0x9e3779b97f4a7c15ull); z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull; z = (z ^ (z >> 27)) * 0x94d049bb133111ebull; return z ^ (z >> 31); }

static void put32(uint8_t *p, uint32_t v) { p[0] = (uint8_t)v; p[1] = (uint8_t)(v >> 8); p[2] = (uint8_t)(v >> 16); p[3] = (uint8_t)(v >> 24); }

static void *rec_worker(void *pp) {
    rec_job_t *j = (rec_job_t *)pp;
    static const uint8_t nib[4] = {1, 2, 4, 8};
    for (;;) {
        pthread_mutex_lock(&j->mu);
        const int64_t a = j->next; j->next += 64;
        pthread_mutex_unlock(&j->mu);
        if (a >= j->n) break;
        const int64_t b = a + 64 < j->n ? a + 64 : j->n;
        for (int64_t r = a; r < b; ++r) {
            uint8_t *p = j->out + j->rec_off[r];
            const int64_t ql = j->qlen[r], nc = j->ncig[r];
            const int64_t len = 36 + 11 + 4 * nc + (ql + 1) / 2 + ql;
            put32(p, (uint32_t)(len - 4)); put32(p + 4, (uint32_t)j->tid[r]); put32(p + 8, j->pos[r]);
            p[12] = 11; p[13] = 60; p[14] = (uint8_t)j->bin[r]; p[15] = (uint8_t)(j->bin[r] >> 8);
            p[16] = (uint8_t)nc; p[17] = (uint8_t)(nc >> 8); p[18] = 0; p[19] = 0;
            put32(p + 20, (uint32_t)ql); put32(p + 24, 0xffffffffu); put32(p + 28, 0xffffffffu); put32(p + 32, 0);
            int64_t id = j->first_index + r;
            p[36] = 'r';
            for (int d = 9; d >= 1; --d) { p[36 + d] = (uint8_t)('0' + id % 10); id /= 10; }
            p[46] = 0;
            memcpy(p + 47, j->cigar + j->cig_off[r], (size_t)(4 * nc));
            uint8_t *s = p + 47 + 4 * nc;
            uint64_t h0 = (uint64_t)(j->first_index + r) * 0xd1342543de82ef95ull + j->seed, st = sm64(&h0) ^ (j->seed * 0x2545f4914f6cdd1dull);   /* independent stream per read */
            st = sm64(&st);
            const int64_t sl = (ql + 1) / 2;
            int64_t i = 0;
            for (; i + 8 <= sl; i += 8) { uint64_t x = sm64(&st); for (int k = 0; k < 8; ++k, x >>= 8) s[i + k] = (uint8_t)(nib[x & 3] << 4 | nib[(x >> 2) & 3]); }
            if (i < sl) { uint64_t x = sm64(&st); for (; i < sl; ++i, x >>= 8) s[i] = (uint8_t)(nib[x & 3] << 4 | nib[(x >> 2) & 3]); }
            if (ql & 1) s[sl - 1] &= 0xf0;
            uint8_t *q = s + sl;
            i = 0;
            for (; i + 8 <= ql; i += 8) { uint64_t x = sm64(&st); for (int k = 0; k < 8; ++k, x >>= 8) q[i + k] = (uint8_t)(10 + (((x & 255) * 31) >> 8)); }
            if (i < ql) { uint64_t x = sm64(&st); for (; i < ql; ++i, x >>= 8) q[i] = (uint8_t)(10 + (((x & 255) * 31) >> 8)); }
        }
    }
    return NULL;
}

int bam_records(int64_t n, int64_t first_index, const int32_t *tid, const uint32_t *pos, const uint16_t *bin, const uint32_t *ncig,
                const int64_t *cig_off, const uint32_t *cigar, const int64_t *qlen, const int64_t *rec_off, uint64_t seed, int threads, uint8_t *out) {
    rec_job_t j = {n, first_index, tid, pos, bin, ncig, cig_off, cigar, qlen, rec_off, seed, out, 0, PTHREAD_MUTEX_INITIALIZER};
    if (threads < 1) threads = 1;
    if (threads > 64) threads = 64;
    pthread_t th[64];
    int made = 0;
    for (int t = 0; t < threads - 1; ++t) if (pthread_create(&th[made], NULL, rec_worker, &j) == 0) ++made;
    rec_worker(&j);
    for (int t = 0; t < made; ++t) pthread_join(th[t], NULL);
    return 0;
}
