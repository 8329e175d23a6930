/* TEST INFRASTRUCTURE ONLY.  CPU restatement ("oracle") of SVTrek's `audt` breakpoint-refinement path.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this
 * file's library.  The product (svtrek_b200/csrc) never links, imports or calls it.
 *
 * Parity status: PINNED.  tests/golden/ holds vectors captured from the unmodified reference compiled from
 * /root/reference (oracle/build_ref.sh -> oracle/_ref/libsvtrek_ref.so and oracle/_ref/svtrek; written by
 * tools/make_golden.py, which asserts oracle == reference on every vector as it goes), and
 * tests/test_oracle_golden.py checks every function below against them wherever the tests run.  Where the
 * compiled reference is present, tests/test_golden_drift.py regenerates the vectors and fails on any drift.
 *
 * The algorithm is restated from the reference's behaviour, not copied: reads live in the SoA layout the
 * GPU path uses (no HTSlib), selection is a range filter instead of an index iterator, the consensus is a
 * straight transcription of the rules in SURVEY.md Appendix A.  Each function cites the reference lines it
 * restates.  Compile with -fwrapv: the reference's int arithmetic is reproduced with wrapping semantics. */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define OP_INS 1
#define OP_DEL 2
#define OP_SOFT 4
#define SV_MIN_LENGTH 50 /* params.h:33 (__SV_MIN_LENGTH) */

enum { KIND_DEL_START = 0, KIND_DEL_END = 1, KIND_INS = 2, KIND_POINT = 3 };

typedef struct {
    int32_t n_tid;
    const int64_t *tid_off;    /* [n_tid+1] */
    const uint32_t *pos;       /* bam1_core_t.pos (0-based) */
    const uint32_t *end;       /* bam_endpos() */
    const uint32_t *n_cigar;
    const uint64_t *cigar_off; /* word offset */
    const uint32_t *cigar;     /* len<<4 | op */
} orc_reads;

typedef struct {
    int wider, median, narrow, range, interval, min_count; /* params.h:27-32 defaults 20000,10000,2000,500,5,3 */
} orc_params;

static int iabs(int a) { return a < 0 ? -a : a; } /* refinement.h:7 macro, on int */

static int cmp_int(const void *a, const void *b) {
    int x = *(const int *)a, y = *(const int *)b;
    return (x > y) - (x < y);
}

/* refinement.c:3-10 */
int orc_lower_bound(const int *a, int n, int x) {
    for (int i = 0; i < n; i++)
        if (a[i] > x) return i == 0 ? 0 : i - 1;
    return n - 1;
}
/* refinement.c:12-19 (note: first element *smaller* than x -- on ascending input this is 0 or n-1) */
int orc_upper_bound(const int *a, int n, int x) {
    for (int i = 0; i < n; i++)
        if (a[i] < x) return i;
    return n - 1;
}

/* refinement.c:41-101.  loc is sorted in place (utils.c:4-25 is a plain ascending sort of ints).
 * n == 0 with min_count <= 0 reads out of bounds in the reference; defined here as -1. */
int orc_consensus_pos(int *loc, int n, int pos, int min_count, int interval, int range) {
    if (n < min_count) return -1;
    if (n <= 0) return -1;
    qsort(loc, (size_t)n, sizeof(int), cmp_int);
    int best_l = -1, maxc_l = min_count - 1, dist_l = 0x7fffffff;
    int best_r = -1, maxc_r = min_count - 1, dist_r = 0x7fffffff;

    int point = orc_lower_bound(loc, n, pos + SV_MIN_LENGTH / 2);
    for (int i = point; 0 <= i && iabs(pos - loc[i]) < range; i--) {
        int count = 1;
        uint64_t total = (uint64_t)(int64_t)loc[i];
        for (int j = i - 1; 0 <= j && loc[i] <= loc[j] + interval; j--) { count++; total += (uint64_t)(int64_t)loc[j]; }
        int cand = (int)((total + (uint64_t)(int64_t)(count / 2)) / (uint64_t)(int64_t)count);
        if (count > maxc_l) {
            if (iabs(pos - cand) < interval) return cand;
            else if (iabs(pos - cand) < dist_l) { maxc_l = count; best_l = cand; dist_l = iabs(pos - cand); }
        }
    }
    point = orc_upper_bound(loc, n, pos - SV_MIN_LENGTH / 2);
    for (int i = point; i < n && iabs(pos - loc[i]) < range; i++) {
        int count = 1;
        uint64_t total = (uint64_t)(int64_t)loc[i];
        for (int j = i + 1; j < n && loc[j] <= loc[i] + interval; j++) { count++; total += (uint64_t)(int64_t)loc[j]; }
        int cand = (int)((total + (uint64_t)(int64_t)(count / 2)) / (uint64_t)(int64_t)count);
        if (count > maxc_r) {
            if (iabs(pos - cand) < interval) return cand;
            else if (iabs(pos - cand) < dist_r) { maxc_r = count; best_r = cand; dist_r = iabs(pos - cand); }
        }
    }
    return dist_l < dist_r ? best_l : best_r;
}

typedef struct { int *v; int n, cap; } ivec;
static void push(ivec *s, int x) {
    if (s->n == s->cap) { s->cap = s->cap ? s->cap * 2 : 128; s->v = (int *)realloc(s->v, sizeof(int) * (size_t)s->cap); }
    s->v[s->n++] = x;
}

/* Which reads a window sees: sam_itr_queryi(idx, chrom-1, start-1, end-1) (refinement.c:114) with the
 * arguments formed in uint32_t; hts_itr_query returns NULL when end < beg or tid is out of range
 * (hts.c:3081-3090); hts_itr_next yields records with pos < end && endpos > beg, stopping at the first
 * pos >= end (hts.c:3905-3921).  Then the CIGAR walk of refinement.c:117-160 / 183-222 / 245-271 / 293-320. */
static void walk_window(const orc_reads *R, int kind, int chrom, uint32_t w_start, uint32_t w_end, ivec *sig) {
    int tid = chrom - 1;
    if (tid < 0 || tid >= R->n_tid) return;
    int64_t beg = (int64_t)(uint32_t)(w_start - 1u), endq = (int64_t)(uint32_t)(w_end - 1u);
    if (endq < beg) return;
    for (int64_t r = R->tid_off[tid]; r < R->tid_off[tid + 1]; r++) {
        int64_t rpos = (int64_t)R->pos[r];
        if (rpos >= endq) break;
        if (!((int64_t)R->end[r] > beg)) continue;
        uint32_t n = R->n_cigar[r];
        if (n == 0) continue; /* reference reads cigar[-1] / cigar[0] here (UB); defined as "no signature" */
        const uint32_t *cg = R->cigar + R->cigar_off[r];
        uint32_t ref = R->pos[r];
        int stopped = 0;
        for (uint32_t i = 0; i < n; i++) {
            uint32_t op = cg[i] & 0xf, len = cg[i] >> 4;
            if (kind == KIND_DEL_START && op == OP_DEL && SV_MIN_LENGTH < len) push(sig, (int)ref);
            else if (kind == KIND_DEL_END && op == OP_DEL && SV_MIN_LENGTH < len) push(sig, (int)(ref + len + 1u));
            else if (kind == KIND_INS && op == OP_INS && SV_MIN_LENGTH <= len) push(sig, (int)ref);
            /* KIND_POINT: refine_point is only ever called with SV_INV and tests sv_type == SV_INS: never emits */
            if (op != OP_INS && op != OP_SOFT) ref += len;
            if (ref > w_end) { stopped = 1; break; }
        }
        if (kind == KIND_DEL_START) { /* trailing soft clip, only when the walk ran to the end (refinement.c:120,142,147) */
            if (!stopped && (cg[n - 1] & 0xf) == OP_SOFT && w_start <= ref && ref <= w_end) push(sig, (int)ref);
        } else if (kind == KIND_DEL_END) { /* leading soft clip; records the POST-walk cursor + 1 (refinement.c:210-220) */
            if ((cg[0] & 0xf) == OP_SOFT && w_start <= R->pos[r] && R->pos[r] <= w_end) push(sig, (int)(ref + 1u));
        }
    }
}

/* refine_start / refine_end / refine_ins / refine_point.  sig_out (optional) receives the raw multiset. */
int orc_refine(const orc_reads *R, int kind, int chrom, uint32_t w_start, uint32_t w_end, uint32_t pos,
               int min_count, int interval, int range, int *sig_out, int sig_cap, int *n_sig) {
    ivec s = {0, 0, 0};
    walk_window(R, kind, chrom, w_start, w_end, &s);
    if (n_sig) *n_sig = s.n;
    if (sig_out) memcpy(sig_out, s.v, sizeof(int) * (size_t)(s.n < sig_cap ? s.n : sig_cap));
    int r = orc_consensus_pos(s.v, s.n, (int)pos, min_count, interval, range);
    free(s.v);
    return r;
}

/* sliding_window.c:8-97.  Per-tile lines go to `out` when non-NULL (format of sliding_window.c:87). */
int orc_sliding_window_ins(const orc_reads *R, int chrom, uint32_t start, uint32_t end, int min_count,
                           int windowSize, int slideSize, FILE *out) {
    int best_all = -1, max_all = 0;
    for (uint32_t sub_start = start; sub_start < end; sub_start += (uint32_t)windowSize) {
        uint32_t sub_end = sub_start + (uint32_t)windowSize;
        if (sub_end > end) sub_end = end;
        ivec s = {0, 0, 0};
        walk_window(R, KIND_INS, chrom, sub_start, sub_end, &s);
        if (s.n == 0) { free(s.v); continue; }
        qsort(s.v, (size_t)s.n, sizeof(int), cmp_int);
        int best = -1, max_support = 0;
        for (int i = 0; i < s.n; i += slideSize) {
            int e = i;
            while (e < s.n && (s.v[e] - s.v[i]) <= windowSize) e++;
            int support = e - i;
            if (support >= min_count && support > max_support) {
                max_support = support;
                int sum = 0;
                for (int j = i; j < e; j++) sum += s.v[j];
                best = (sum + support / 2) / support;
            }
            if (slideSize <= 0) break; /* reference would loop forever */
        }
        if (best != -1) {
            if (out) fprintf(out, "INS Discovery in window [%d, %d] at position %d with support %d\n", sub_start, sub_end, best, max_support);
            if (max_support > max_all) { max_all = max_support; best_all = best; }
        }
        free(s.v);
        if (windowSize <= 0) break;
    }
    return best_all;
}

/* same, per-tile lines appended to the file at out_path (ctypes-friendly) */
int orc_sliding_window_ins_path(const orc_reads *R, int chrom, uint32_t start, uint32_t end, int min_count,
                                int windowSize, int slideSize, const char *out_path) {
    FILE *f = out_path ? fopen(out_path, "a") : NULL;
    int r = orc_sliding_window_ins(R, chrom, start, end, min_count, windowSize, slideSize, f);
    if (f) fclose(f);
    return r;
}

/* ---------------------------------------------------------------------------------------------------
 * VCF record parsing + window construction + output formatting: audit.c:50-248 (thread_func).        */
enum { SV_UNKNOWN = 0, SV_INS, SV_DEL, SV_INV, SV_DUP, SV_TRA, SV_BND };

static int sv_from_str(const char *s) { /* audit.c:3-11 */
    if (!strcmp(s, "INS") || !strcmp(s, "INS:ME")) return SV_INS;
    if (!strcmp(s, "DEL") || !strcmp(s, "DEL:ME")) return SV_DEL;
    if (!strcmp(s, "INV")) return SV_INV;
    if (!strcmp(s, "DUP")) return SV_DUP;
    if (!strcmp(s, "TRA")) return SV_TRA;
    if (!strcmp(s, "BND")) return SV_BND;
    return SV_UNKNOWN;
}

typedef struct { int type, chrom; uint32_t pos, end; } orc_record;

/* returns 1 = record, 0 = skipped (silently or with a message on err), -1 = "[ERROR] Unkown type." */
int orc_parse_line(char *line, orc_record *rec, FILE *err) {
    char *save, *asave;
    char *chrom = strtok_r(line, "\t", &save);
    char *index = strtok_r(NULL, "\t", &save);
    if (index == NULL) { if (err) fprintf(err, "VCF: no index at line: %s\n", line); return 0; } /* audit.c:67-71 */
    strtok_r(NULL, "\t", &save);
    char *seq = strtok_r(NULL, "\t", &save);
    char *alt = strtok_r(NULL, "\t", &save);
    if (!seq || !alt) return 0; /* the reference dereferences NULL here; defined as skipped */
    int seq_len = (int)strlen(seq);
    size_t max_alt = 0, min_alt = 0x7FFFFFFF;
    for (char *t = strtok_r(alt, ",", &asave); t; t = strtok_r(NULL, ",", &asave)) { /* audit.c:79-90 */
        size_t l = strlen(t);
        if (l > max_alt) max_alt = l;
        if (l < min_alt) min_alt = l;
    }
    strtok_r(NULL, "\t", &save);
    strtok_r(NULL, "\t", &save);
    char *info = strtok_r(NULL, "\t", &save);
    if (!info) return 0; /* reference: strstr(NULL, ...) crash */

    rec->chrom = strncmp(chrom, "chr", 3) == 0 ? atoi(chrom + 3) : atoi(chrom); /* audit.c:101-105 */
    rec->pos = (uint32_t)strtol(index, NULL, 10);                               /* audit.c:108 */
    if (rec->pos == 0 && index[0] != '0') { if (err) fprintf(err, "[ERROR] Conversion error to pos %s\n", index); return 0; }

    char *sv = strstr(info, "SVTYPE="); /* audit.c:116-142 */
    if (sv) {
        sv += 7;
        char *e = strchr(sv, ';');
        size_t len = e ? (size_t)(e - sv) : strlen(sv);
        char buf[16] = {0};
        if (len >= sizeof buf) len = sizeof buf - 1;
        memcpy(buf, sv, len);
        rec->type = sv_from_str(buf);
    } else if (seq_len == 1 && SV_MIN_LENGTH < max_alt) rec->type = SV_INS;
    else if (SV_MIN_LENGTH < seq_len && min_alt == 1) rec->type = SV_DEL;
    else return 0;

    char *es = strstr(info, "END="); /* audit.c:145-166: first match, may sit inside CIEND= */
    if (es) {
        es += 4;
        char *e = strchr(es, ';');
        size_t len = e ? (size_t)(e - es) : strlen(es);
        char buf[32] = {0};
        if (len >= sizeof buf) len = sizeof buf - 1;
        memcpy(buf, es, len);
        rec->end = (uint32_t)strtol(buf, NULL, 10);
        if (rec->end == 0 && buf[0] != '0') return 0;
    } else rec->end = rec->pos + (uint32_t)strlen(seq);

    if ((rec->type == SV_DEL || rec->type == SV_INV) && rec->end - rec->pos < SV_MIN_LENGTH) return 0; /* audit.c:168-173 */
    if (rec->type == SV_INS || rec->type == SV_DEL || rec->type == SV_INV) return 1;
    if (err) fprintf(err, "[ERROR] Unkown type.\n"); /* audit.c:233-235 */
    return -1;
}

/* One record -> its stdout line (audit.c:175-232).  Returns number of lines written (0 or 1). */
int orc_audit_record(const orc_reads *R, const orc_record *rec, const orc_params *P, FILE *out) {
    uint32_t pos = rec->pos, end = rec->end;
    if (rec->type == SV_INS) {
        uint32_t res = (uint32_t)orc_refine(R, KIND_INS, rec->chrom, pos - (uint32_t)P->median, pos + (uint32_t)P->median, pos,
                                            P->min_count, P->interval, P->range, NULL, 0, NULL);
        if (res == 0xFFFFFFFFu) fprintf(out, "(INS) chr: %d, org pos: %u, ref pos: NA\n", rec->chrom, pos);
        else fprintf(out, "(INS) chr: %d, org pos: %u, ref pos: %u, diff: %d\n", rec->chrom, pos, res, (int)(res - pos));
        return 1;
    }
    if (rec->type == SV_DEL) {
        if (!(SV_MIN_LENGTH < end - pos)) return 0;
        uint32_t rs = (uint32_t)orc_refine(R, KIND_DEL_START, rec->chrom, pos - (uint32_t)P->wider, pos + (uint32_t)P->narrow, pos,
                                           P->min_count, P->interval, P->range, NULL, 0, NULL);
        uint32_t re = (uint32_t)orc_refine(R, KIND_DEL_END, rec->chrom, end - (uint32_t)P->narrow, end + (uint32_t)P->narrow, end,
                                           P->min_count, P->interval, P->range, NULL, 0, NULL);
        fprintf(out, "(DEL) chr: %d, org pos: %u, org end: %u, ref pos: ", rec->chrom, pos, end);
        if (rs == 0xFFFFFFFFu) fprintf(out, "NA, ref end: "); else fprintf(out, "%d, ref end: ", (int)rs);
        if (re == 0xFFFFFFFFu) fprintf(out, "NA, "); else fprintf(out, "%d, ", (int)re);
        if (rs == 0xFFFFFFFFu) fprintf(out, "diff pos: NA, "); else fprintf(out, "diff pos: %d, ", (int)(rs - pos));
        if (re == 0xFFFFFFFFu) fprintf(out, "diff end: NA\n"); else fprintf(out, "diff end: %d\n", (int)(re - end));
        return 1;
    }
    if (rec->type == SV_INV) {
        if (!(SV_MIN_LENGTH < end - pos)) return 0;
        uint32_t rs = (uint32_t)orc_refine(R, KIND_POINT, rec->chrom, pos - (uint32_t)P->wider, pos + (uint32_t)P->wider, pos,
                                           P->min_count, P->interval, P->range, NULL, 0, NULL);
        uint32_t re = (uint32_t)orc_refine(R, KIND_POINT, rec->chrom, end - (uint32_t)P->wider, end + (uint32_t)P->wider, end,
                                           P->min_count, P->interval, P->range, NULL, 0, NULL);
        fprintf(out, "(INV) chr: %d, org pos: %u, org end: %u, ref pos: %u, ref end: %u\n", rec->chrom, pos, end, rs, re);
        return 1;
    }
    return 0;
}

/* process_vcf (audit.c:250-357) at -t 1 with NO tail-drop (every line is processed): reads the
 * VCF file, writes the reference's stdout to out_path and its stderr messages to err_path (may be NULL).
 * Returns the number of record lines written, or -1 on I/O error. */
long orc_audit_file(const orc_reads *R, const char *vcf_path, const orc_params *P, const char *out_path, const char *err_path) {
    FILE *vf = fopen(vcf_path, "r");
    FILE *out = fopen(out_path, "w");
    FILE *err = err_path ? fopen(err_path, "w") : NULL;
    if (!vf || !out) return -1;
    fprintf(out, "[INFO] Started processing variation file.\n");
    char *line = NULL;
    size_t cap = 0;
    ssize_t len;
    long n = 0;
    while ((len = getline(&line, &cap, vf)) >= 0) {
        if (len < 2 || line[0] == '#') continue; /* audit.c:324 */
        if (line[len - 1] == '\n') line[--len] = 0;
        orc_record rec;
        if (orc_parse_line(line, &rec, err) == 1) n += orc_audit_record(R, &rec, P, out);
    }
    fprintf(out, "[INFO] Ended processing variation file\n");
    free(line);
    fclose(vf); fclose(out);
    if (err) fclose(err);
    return n;
}

/* Batched entry for timing / parity of the device batch seam: window arrays in, results out. */
void orc_refine_batch(const orc_reads *R, int n_win, const int32_t *chrom, const uint32_t *w_start, const uint32_t *w_end,
                      const uint32_t *pos, const uint8_t *kind, int min_count, int interval, int range, int32_t *result) {
    for (int i = 0; i < n_win; i++)
        result[i] = orc_refine(R, kind[i], chrom[i], w_start[i], w_end[i], pos[i], min_count, interval, range, NULL, 0, NULL);
}
