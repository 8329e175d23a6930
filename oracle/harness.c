/* TEST INFRASTRUCTURE ONLY -- direct-call shim over the UNMODIFIED reference objects
 * (refinement.o, sliding_window.o, utils.o built from /root/reference by oracle/build_ref.sh).
 * It exposes the reference's non-static hot-path functions to pytest through ctypes and lets the
 * tests observe the raw signature multiset of a window: the harness copies of refinement.c and
 * sliding_window.c are compiled with -Dquicksort=svtk_hook_quicksort, and quicksort() is called
 * exactly once per window with the unsorted signature array (refinement.c:54, sliding_window.c:65).
 * No reference source is reproduced here. */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <stdint.h>
#include "params.h"

/* reference symbols (refinement.c / sliding_window.c / utils.c) */
int consensus_pos(int *locations, int size, int pos, int consensus_min_count, int consensus_interval, int consensus_interval_range);
int lower_bound(int *arr, int size, int location);
int upper_bound(int *arr, int size, int location);
int refine_start(sv_type_t sv_type, int chrom, interval inter, uint32_t imprecise_pos, t_arg *params);
int refine_end(sv_type_t sv_type, int chrom, interval inter, uint32_t imprecise_pos, t_arg *params);
int refine_point(sv_type_t sv_type, int chrom, interval inter, uint32_t imprecise_pos, t_arg *params);
int refine_ins(int chrom, interval inter, uint32_t imprecise_pos, t_arg *params);
int sliding_window_ins(int chrom, interval inter, t_arg *params, int windowSize, int slideSize);
void quicksort(int *array, int low, int high);

static int *g_cap = NULL;
static int g_cap_n = 0, g_cap_alloc = 0, g_cap_on = 0;

void svtk_hook_quicksort(int *array, int low, int high) {
    if (g_cap_on && low == 0) {
        int n = high + 1;
        if (n > g_cap_alloc) { g_cap = (int *)realloc(g_cap, sizeof(int) * (size_t)n); g_cap_alloc = n; }
        if (n > 0) memcpy(g_cap, array, sizeof(int) * (size_t)n);
        g_cap_n = n < 0 ? 0 : n;
        g_cap_on = 0; /* only the top-level call carries the whole array */
    }
    quicksort(array, low, high);
}

int ref_consensus_pos(const int *arr, int n, int pos, int minc, int interval, int range) {
    int *tmp = (int *)malloc(sizeof(int) * (size_t)(n > 0 ? n : 1));
    if (n > 0) memcpy(tmp, arr, sizeof(int) * (size_t)n);
    g_cap_on = 0;
    int r = consensus_pos(tmp, n, pos, minc, interval, range);
    free(tmp);
    return r;
}
int ref_lower_bound(int *arr, int n, int x) { return lower_bound(arr, n, x); }
int ref_upper_bound(int *arr, int n, int x) { return upper_bound(arr, n, x); }

typedef struct { t_arg t; } ref_handle;

void *ref_open(const char *bam) {
    ref_handle *h = (ref_handle *)calloc(1, sizeof(ref_handle));
    h->t.hargs.fp_in = hts_open(bam, "r");
    if (!h->t.hargs.fp_in) { free(h); return NULL; }
    h->t.hargs.bam_hdr = sam_hdr_read(h->t.hargs.fp_in);
    h->t.hargs.bam_file_index = sam_index_load(h->t.hargs.fp_in, bam);
    if (!h->t.hargs.bam_hdr || !h->t.hargs.bam_file_index) return NULL;
    return h;
}
void ref_close(void *hv) {
    ref_handle *h = (ref_handle *)hv;
    if (!h) return;
    hts_idx_destroy(h->t.hargs.bam_file_index);
    bam_hdr_destroy(h->t.hargs.bam_hdr);
    sam_close(h->t.hargs.fp_in);
    free(h);
}

/* kind: 0 = refine_start(SV_DEL), 1 = refine_end(SV_DEL), 2 = refine_ins, 3 = refine_point(SV_INV).
 * Returns the reference's result. If sig_out != NULL the raw signature multiset (unsorted, reference
 * emission order) is copied there (up to sig_cap) and *n_sig receives its size. */
int ref_refine(void *hv, int kind, int chrom, uint32_t start, uint32_t end, uint32_t pos,
               int minc, int interval_, int range, int *sig_out, int sig_cap, int *n_sig) {
    ref_handle *h = (ref_handle *)hv;
    interval w = { start, end };
    t_arg *t = &h->t;
    int r = -1;
    for (int pass = (sig_out || n_sig) ? 0 : 1; pass < 2; pass++) {
        if (pass == 0) { t->consensus_min_count = 1; g_cap_on = 1; g_cap_n = 0; }
        else           { t->consensus_min_count = minc; g_cap_on = 0; }
        t->consensus_interval = interval_;
        t->consensus_interval_range = range;
        switch (kind) {
        case 0: r = refine_start(SV_DEL, chrom, w, pos, t); break;
        case 1: r = refine_end(SV_DEL, chrom, w, pos, t); break;
        case 2: r = refine_ins(chrom, w, pos, t); break;
        default: r = refine_point(SV_INV, chrom, w, pos, t); break;
        }
        if (pass == 0) {
            g_cap_on = 0;
            if (n_sig) *n_sig = g_cap_n;
            if (sig_out) memcpy(sig_out, g_cap, sizeof(int) * (size_t)(g_cap_n < sig_cap ? g_cap_n : sig_cap));
        }
    }
    return r;
}

int ref_sliding_window_ins(void *hv, int chrom, uint32_t start, uint32_t end, int minc, int windowSize, int slideSize) {
    ref_handle *h = (ref_handle *)hv;
    interval w = { start, end };
    h->t.consensus_min_count = minc;
    g_cap_on = 0;
    fflush(stdout);
    int r = sliding_window_ins(chrom, w, &h->t, windowSize, slideSize);
    fflush(stdout);
    return r;
}
