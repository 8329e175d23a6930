/* svtk.h -- C ABI of libsvtk_b200.so: the B200 (sm_100a) drop-in for SVTrek's `audt` breakpoint-refinement
 * hot path.  Plain C, plain pointers and sizes; no CUDA or torch types cross this boundary.
 *
 * The reference (akmami/SVTrek) has no FFI of its own; the seams this library replaces are
 *   - the per-record seam   deletion()/insertion()/inversion()           refinement.h:25,42,60
 *     (called only from thread_func, audit.c:180,195,228), which fan out to
 *     refine_start/refine_end/refine_ins/refine_point                    refinement.c:103,169,278,231
 *     and consensus_pos                                                  refinement.c:41
 *   - the batch seam        process_vcf(audt_args*)                      audit.c:250
 *   - the (dead) discovery entry sliding_window_ins()                    sliding_window.c:8
 * and the vocabulary is params.h's: `interval` (params.h:123-126), the six tunables of audt_args/t_arg
 * (params.h:57-62, 82-87), CIGAR op codes (params.h:10-18), __SV_MIN_LENGTH (params.h:33).
 *
 * Where the reference hands each call an open BAM (t_arg.hargs, params.h:75-79) and re-inflates it per
 * window, this library holds the decoded reads once, in HBM, as SoA arrays (svtk_reads): the caller (the
 * host driver, after HTSlib / the bundled BGZF reader decoded the BAM) packs pos / endpos / n_cigar /
 * CIGAR words in coordinate order and uploads them with svtk_load_reads().  "Which reads does a window
 * see" (sam_itr_queryi + sam_itr_next, hts.c:3049-3090, 3905-3921) is reproduced on the device.
 *
 * Conventions: every function returns 0 on success and a negative svtk_status on failure; the message is
 * available from svtk_last_error().  Results use the reference's own encoding: -1 (0xFFFFFFFF) == "NA".
 * The caller owns every host buffer it passes.  One ctx drives one GPU; a ctx is not re-entrant (use one
 * per host thread / per GPU, like the reference's one t_arg per worker, audit.c:269-285).
 * There is no CPU fallback: without a CUDA device every compute entry point fails with SVTK_ERR_CUDA. */
#ifndef SVTK_H
#define SVTK_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SVTK_VERSION 100 /* 0.1.0 */

typedef enum {
    SVTK_OK = 0,
    SVTK_ERR_ARG = -1,    /* bad argument */
    SVTK_ERR_CUDA = -2,   /* CUDA runtime error or no device */
    SVTK_ERR_ALLOC = -3,  /* host or device allocation failed */
    SVTK_ERR_STATE = -4,  /* e.g. no reads loaded */
    SVTK_ERR_IO = -5      /* file could not be opened / parsed */
} svtk_status;

/* window kinds: which reference walker a window runs */
enum {
    SVTK_DEL_START = 0, /* refine_start(SV_DEL, ...)  refinement.c:103 */
    SVTK_DEL_END = 1,   /* refine_end(SV_DEL, ...)    refinement.c:169 */
    SVTK_INS = 2,       /* refine_ins(...)            refinement.c:278 */
    SVTK_POINT = 3      /* refine_point(SV_INV, ...)  refinement.c:231 -- never collects a position */
};

/* params.h:123-126 */
typedef struct { uint32_t start, end; } svtk_interval;

/* the six tunables of audt_args / t_arg (params.h:57-62); defaults params.h:27-32 */
typedef struct {
    int32_t wider_interval;           /* 20000 */
    int32_t median_interval;          /* 10000 */
    int32_t narrow_interval;          /* 2000  */
    int32_t consensus_interval_range; /* 500   */
    int32_t consensus_interval;       /* 5     */
    int32_t consensus_min_count;      /* 3     */
} svtk_params;

/* Decoded reads, SoA, sorted by (tid, pos) == BAM coordinate order.  cigar words are BAM-native
 * (oplen << 4 | op, htslib/sam.h:124-125).  `end` is bam_endpos() (sam.c:638-643): pos + max(1, reference
 * length), and pos + 1 for reads flagged unmapped.  Reads with n_cigar == 0 contribute no signature
 * (the reference indexes cigar[-1] / cigar[0] for them, refinement.c:120,210 -- undefined behaviour). */
typedef struct {
    int32_t n_tid;
    int64_t n_reads;
    int64_t n_words;
    const int64_t *tid_off;    /* [n_tid + 1] read index range of each contig */
    const uint32_t *pos;       /* [n_reads] bam1_core_t.pos, 0-based */
    const uint32_t *end;       /* [n_reads] */
    const uint32_t *n_cigar;   /* [n_reads] */
    const uint64_t *cigar_off; /* [n_reads] word offset of each read's CIGAR inside `cigar`; CIGARs are packed back to
                                  back in read order: cigar_off[0] == 0, cigar_off[r] + n_cigar[r] == cigar_off[r+1] */
    const uint32_t *cigar;     /* [n_words] */
} svtk_reads;

/* A batch of windows, SoA.  One window == one refine_* call of the reference:
 * (kind, chrom, inter, imprecise_pos) -- refinement.c:103,169,231,278.  chrom is 1-based (tid = chrom-1). */
typedef struct {
    int32_t n_win;
    const int32_t *chrom;
    const uint32_t *start; /* inter.start */
    const uint32_t *end;   /* inter.end   */
    const uint32_t *pos;   /* imprecise_pos */
    const uint8_t *kind;   /* SVTK_DEL_START .. SVTK_POINT */
} svtk_windows;

/* Work counters of the last instrumented pass (svtk_measure_windows): the algorithmic bytes of SURVEY 8(d). */
typedef struct {
    uint64_t n_windows;
    uint64_t n_pairs;          /* (window, read) pairs the reference's iterators would yield */
    uint64_t n_words_visited;  /* CIGAR words the reference's walk loops touch */
    uint64_t n_signatures;     /* positions pushed into the per-window arrays */
    uint64_t n_overflow;       /* windows that left the shared-memory fast path */
    uint64_t algorithmic_bytes;/* 4*words + 16*pairs + 4*signatures + 24*windows + 4*windows(result) */
} svtk_stats;

/* Live timing of the dominant kernel (the fused CIGAR-walk + consensus kernel), taken with CUDA events on the
 * ctx stream while svtk_set_profiling(ctx, 1) is in force; plus the bytes the host-buffer entry points moved. */
typedef struct {
    double refine_kernel_ms;          /* per-window kernel: sum of launch durations since profiling was switched on */
    uint64_t refine_kernel_launches;
    double index_kernel_ms;           /* per-read CIGAR pass (signature index) */
    uint64_t index_kernel_launches;
    uint64_t h2d_bytes, d2h_bytes;    /* since svtk_create */
} svtk_timing;

/* Work counters of the product path (svtk_measure_index): the bytes each of its two kernels has to move. */
typedef struct {
    uint64_t n_reads, n_words, n_chunks;   /* the resident read set; chunks of 32 reads */
    uint64_t n_events, n_segments;         /* signature events (D > 50, I >= 50, soft clips) the index holds */
    uint64_t index_bytes;                  /* one index pass: 4*n_words + 24*n_reads + 16*n_events + 16*n_chunks */
    uint64_t n_windows;
    uint64_t refine_events_scanned, refine_segments, refine_reads_in_range;
    uint64_t n_signatures, n_overflow;
    uint64_t refine_bytes;                 /* one refine pass: 16*events scanned + 16*segments + 28*n_windows */
} svtk_index_stats;

typedef struct svtk_ctx svtk_ctx;

int svtk_version(void);
const char *svtk_last_error(void);
int svtk_device_count(void);

/* lifetime.  `cuda_stream` may be NULL (the ctx creates its own non-blocking stream) or a cudaStream_t
 * owned by the caller (so a harness can bracket launches with its own events). */
svtk_ctx *svtk_create(int device, const svtk_params *params, void *cuda_stream);
void svtk_destroy(svtk_ctx *ctx);
int svtk_set_params(svtk_ctx *ctx, const svtk_params *params);
int svtk_get_params(const svtk_ctx *ctx, svtk_params *params);

/* reads: host SoA -> HBM (pinned staging, async copies on the ctx stream).  Replaces what
 * hts_open/sam_index_load (audit.c:270-272) give the reference: access to every read of the BAM. */
int svtk_load_reads(svtk_ctx *ctx, const svtk_reads *host_reads);
/* adopt arrays that already live in device memory (pointers in *dev_reads are device pointers; tid_off is a
 * HOST pointer).  The ctx does not take ownership.  Used when the packer itself runs on the GPU. */
int svtk_attach_reads_device(svtk_ctx *ctx, const svtk_reads *dev_reads);

/* ---- BAM in: the whole coordinate-sorted, BAI-indexed BAM -> resident read set, decoded on the GPU.
 * Replaces hts_open + sam_index_load (audit.c:270-272) and everything sam_itr_queryi / sam_itr_next do per window
 * (refinement.c:109-117; bgzf.c:988 bgzf_read_block -> bgzf.c:746 inflate; sam.c:754-795 bam_read1 incl. the CG-tag CIGAR
 * fix-up sam.c:645; sam.c:638 bam_endpos): the file is copied to HBM compressed, every BGZF block is inflated there, the
 * record chain is cut at the record-aligned offsets the .bai holds, and pos / endpos / CIGAR are extracted into the same
 * arrays svtk_load_reads fills.  Every record with refID >= 0 becomes a read (no FLAG / MAPQ filter, like the reference).
 * bai_path NULL: "<bam_path>.bai", else "<bam_path minus .bam>.bai".  Errors (SVTK_ERR_IO): unreadable / corrupt file,
 * stale index, records out of coordinate order, CRC32 mismatch of a block (checked on the GPU, like bgzf.c:776). */
typedef struct {
    uint64_t file_bytes, n_blocks, inflated_bytes;   /* n_blocks / inflated_bytes: of the part of the file that was loaded */
    uint64_t n_reads, n_words;
    uint64_t image_bytes;            /* compressed bytes shipped to the GPU (== file_bytes for a whole-file load) */
    int32_t n_ref;
    uint32_t n_pieces;               /* pieces the record chain was cut into */
    uint32_t lanes_per_warp;         /* BGZF blocks decoded side by side in one warp (1..32, grows with the file) */
    float host_scan_ms;              /* (unused since the block table is built chunk by chunk while the file uploads) */
    float upload_ms;                 /* first to last host -> device copy of the file's bytes */
    float inflate_ms;                /* first inflate launch to the end of the last one (overlaps the upload) */
    float parse_ms, total_ms;        /* CRC + record kernels; wall clock of the whole call */
    float load_span_ms;              /* first upload to the end of the last inflate: what upload + inflate cost together */
    uint32_t n_spans;                /* disjoint byte ranges of the file that were shipped */
    uint32_t inflate_launches;
} svtk_bam_stats;
int svtk_load_bam(svtk_ctx *ctx, const char *bam_path, const char *bai_path, svtk_bam_stats *stats);
/* The same for the reads of references tid_lo .. tid_hi only (region shard of a multi-GPU run, SURVEY 8e: one ctx per GPU, each
 * loads its own contiguous group of references): the index's bin chunks say where those references' records start and end in
 * the file, so only that byte range is shipped and inflated.  Read indices and tid_off are those of the shard (other
 * references are empty); windows on other references refine to NA. */
int svtk_load_bam_refs(svtk_ctx *ctx, const char *bam_path, const char *bai_path, int32_t tid_lo, int32_t tid_hi, svtk_bam_stats *stats);
/* Only what a batch of windows can see (host arrays; chrom / start / end are read): for each window the region its iterator
 * would query -- sam_itr_queryi(idx, chrom - 1, start - 1, end - 1), refinement.c:114 -- is looked up in the .bai the way
 * hts_itr_query does (hts.c:3049-3090: the chunks of every overlapping bin, clipped by the linear index), the chunk lists are
 * merged, and only those BGZF blocks are shipped and inflated.  Region shards of a single-reference job, and sparse VCFs against
 * large BAMs, load a fraction of the file.  Reads beyond what the windows can see are simply not resident. */
int svtk_load_bam_windows(svtk_ctx *ctx, const char *bam_path, const char *bai_path, const svtk_windows *host_windows, svtk_bam_stats *stats);
/* the BAM header's reference dictionary (bam_hdr_t target_name / target_len) after svtk_load_bam */
int svtk_bam_n_ref(const svtk_ctx *ctx);
int svtk_bam_ref(const svtk_ctx *ctx, int tid, const char **name, uint32_t *len);
/* the resident read set copied back to host arrays (any array may be NULL; call once for the sizes).  Parity aid. */
int svtk_copy_reads(svtk_ctx *ctx, int64_t *n_tid, int64_t *n_reads, int64_t *n_words, int64_t *tid_off, uint32_t *pos, uint32_t *end,
                    uint32_t *n_cigar, uint64_t *cigar_off, uint32_t *cigar);

/* ---- inserted sequences behind the refined insertion breakpoints (no call site in the reference, which never
 * looks at the bases: refinement.c:299-310 pushes positions only, and abPOA is linked but never called -- SURVEY F5).
 * For every SVTK_INS window w with refined[w] >= 0 (the result svtk_refine_windows gave for it): every `I` op with
 * len >= min_len (the reference's signature threshold is 50, refinement.c:299) that refine_ins would have pushed for this
 * window -- same reads (sam_itr_queryi selection), same walk and stop rule -- and whose reference position lies within
 * `tol` bases of refined[w] contributes its inserted bases, taken from the 4-bit SEQ of the BAM image svtk_load_bam left in
 * HBM (bam_get_seq / bam_seqi).  One read set per window that has at least one such op, sequences in file order (read, then
 * position in the read), in svtk_poa_consensus' layout: pass {n_sets, set_off, seq_off, bases} on as a svtk_poa_batch.
 * The arrays belong to the ctx and stay valid until the next svtk_ins_sequences / svtk_destroy.  Needs reads loaded with
 * svtk_load_bam (SVTK_ERR_STATE otherwise).  Reads stored without SEQ (`*`) contribute nothing. */
typedef struct {
    int32_t n_sets;
    int64_t n_seqs;
    const int64_t *set_off;   /* [n_sets + 1] */
    const int64_t *seq_off;   /* [n_seqs + 1] */
    const uint8_t *bases;     /* abPOA codes A0 C1 G2 T3 other 4 */
    const int32_t *window;    /* [n_sets]  window index of each set */
    const int32_t *read;      /* [n_seqs]  resident read index of each sequence */
    const uint32_t *ref_pos;  /* [n_seqs]  reference cursor of the I op (what refine_ins pushed) */
} svtk_ins_sets;
int svtk_ins_sequences(svtk_ctx *ctx, const svtk_windows *host_windows, const int32_t *refined, int32_t tol, int32_t min_len, svtk_ins_sets *out);

/* ---- batch seam: N refine_* calls in one go.  result[i] == the reference's return value (-1 == NA). */
/* host buffers in, host buffer out; H2D + kernels + D2H, synchronous */
int svtk_refine_windows(svtk_ctx *ctx, const svtk_windows *host_windows, int32_t *host_result);
/* device buffers in/out; asynchronous on the ctx stream (svtk_sync to wait) */
int svtk_refine_windows_device(svtk_ctx *ctx, const svtk_windows *dev_windows, int32_t *dev_result);
/* same results through the reference-shaped walker (one CIGAR walk per (window, read) pair, like refinement.c) with
 * the work counters of SURVEY 8(d); slower -- an independent cross-check and the reference-equivalent byte count */
int svtk_measure_windows(svtk_ctx *ctx, const svtk_windows *dev_windows, int32_t *dev_result, svtk_stats *stats);
/* One whole pass of the path -- rebuild the signature index of the resident reads, then refine every window -- queued on the
 * ctx stream without waiting for anything: the caller may queue many passes back to back (bench.py's timed steps; a driver that
 * audits several VCFs against one BAM).  svtk_sync waits for all of them and finishes, synchronously, the windows of any pass that
 * left the fast path.  dev_windows / dev_result must stay valid until then. */
int svtk_audit_device(svtk_ctx *ctx, const svtk_windows *dev_windows, int32_t *dev_result);
int svtk_sync(svtk_ctx *ctx);
/* number of kernel launches this ctx has issued so far */
uint64_t svtk_launch_count(const svtk_ctx *ctx);
/* bracket every index / refinement kernel launch with CUDA events (on = 1), or every on-th launch of each (on > 1:
 * a timing event pair serialises the stream around the kernel, ~4 us each on B200), and read the running totals;
 * switching profiling on or off resets them.  No reference counterpart: measurement aid for bench.py's roofline. */
int svtk_set_profiling(svtk_ctx *ctx, int on);
int svtk_get_timing(svtk_ctx *ctx, svtk_timing *timing);
/* The product path walks every CIGAR once (the "signature index": one coalesced pass over the resident reads) and
 * then answers each window from the events it found.  The first svtk_refine_* call after svtk_load_reads builds the
 * index; svtk_index_reads rebuilds it explicitly (asynchronous, ctx stream) -- bench.py calls it every step so that a
 * timed step always contains the whole job.  svtk_measure_index: rebuild + refine with work counters. */
int svtk_index_reads(svtk_ctx *ctx);
int svtk_measure_index(svtk_ctx *ctx, const svtk_windows *dev_windows, int32_t *dev_result, svtk_index_stats *stats);

/* ---- per-record seam: same meaning as refinement.h:25,42,60 with `params` replaced by the ctx. */
int svtk_deletion(svtk_ctx *ctx, int chrom, svtk_interval begin, svtk_interval end, svtk_interval sv_inter, svtk_interval *res_inter);
int svtk_insertion(svtk_ctx *ctx, int chrom, svtk_interval begin, uint32_t pos, uint32_t *res_start);
int svtk_inversion(svtk_ctx *ctx, int chrom, svtk_interval begin, svtk_interval end, svtk_interval sv_inter, svtk_interval *res_inter);

/* ---- pieces of the path, exposed for parity tests against the reference's non-static functions */
/* consensus_pos (refinement.c:41) on n_seg independent arrays: values[offsets[s] .. offsets[s+1]) with its own
 * imprecise position pos[s]; min_count / interval / range come from the ctx params.  Runs on the device. */
int svtk_consensus_pos_batch(svtk_ctx *ctx, const int32_t *values, const int64_t *offsets, int32_t n_seg,
                             const int32_t *pos, int32_t *result);
/* svtk_consensus_pos_batch on DEVICE arrays, allocating nothing.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  dev_values [n_values] / dev_offsets [n_seg + 1] / dev_pos [n_seg] / dev_result [n_seg] are device pointers the caller
 * owns; the kernel is enqueued on the ctx stream and the call returns without waiting (svtk_sync, or synchronise the stream
 * given to svtk_create, before reading dev_result).  The global sort scratch of segments longer than 8 192 values is a
 * file-scope device array of 2 * SVTK_STATIC_MAX_VALUES ints (8 MiB, held by every GPU that loads the library until exit):
 * n_values > SVTK_STATIC_MAX_VALUES fails with SVTK_ERR_ARG and launches nothing.  No cudaMalloc / cudaFree / allocating
 * library call is made; the first call on a device looks the array's address up (cudaGetSymbolAddress) -- make that call
 * outside a capture.  All calls on one GPU share the array: nothing in it has to start at zero (a segment reads back only what
 * it wrote in the same launch), but calls that overlap on the GPU (other streams, other ctx) race.  A segment that does not
 * lie inside [0, n_values] gets SVTK_STATIC_BAD_SEGMENT and touches no memory. */
#define SVTK_STATIC_MAX_VALUES (1 << 20)
#define SVTK_STATIC_BAD_SEGMENT (-2147483647 - 1)
int svtk_consensus_pos_batch_static(svtk_ctx *ctx, const int32_t *dev_values, const int64_t *dev_offsets, int32_t n_seg, int64_t n_values,
                                    const int32_t *dev_pos, int32_t *dev_result);
/* the raw signature multiset of one window (what the reference collects before quicksort, refinement.c:54),
 * ascending.  *n_sig receives the full count even when it exceeds cap. */
int svtk_window_signatures(svtk_ctx *ctx, int kind, int chrom, svtk_interval inter, int32_t *sig_out, int64_t cap, int64_t *n_sig);
/* sliding_window_ins (sliding_window.c:8).  Per-tile "INS Discovery ..." lines (sliding_window.c:87) are
 * written to lines_buf (NUL-terminated, truncated to cap) when it is non-NULL. */
int svtk_sliding_window_ins(svtk_ctx *ctx, int chrom, svtk_interval inter, int window_size, int slide_size,
                            int32_t *best, char *lines_buf, size_t cap);

/* ---- whole path: process_vcf (audit.c:250).  Parses the VCF exactly like thread_func (audit.c:65-173),
 * refines every record on the GPU and writes the reference's stdout lines in VCF order to out_path
 * (NULL or "-" == stdout); the reference's stderr messages go to stderr.  Every record is processed (the
 * reference's end-of-file tail drop, audit.c:35-41,342, is not reproduced).  *n_records (may be NULL)
 * receives the number of (INS)/(DEL)/(INV) lines written. */
int svtk_process_vcf(svtk_ctx *ctx, const char *vcf_path, const char *out_path, int64_t *n_records);
/* A region shard of process_vcf for multi-GPU runs (replaces the per-worker fan-out of audit.c:269-293): the whole
 * VCF is parsed, but only the records on references tid_lo .. tid_hi (chrom - 1) are audited and written.  One ctx per GPU
 * calls this with its own range (and the reads of that range loaded, svtk_load_bam_refs); the shard outputs concatenated in
 * range order are the lines svtk_process_vcf writes for the whole file.  flags: SVTK_VCF_NO_INFO_LINES leaves out the two
 * [INFO] lines (the merger writes them once); SVTK_VCF_OWNS_UNPLACED makes this shard the one that reports lines no shard
 * can place (malformed lines, references outside the dictionary) -- give it to exactly one shard. */
enum { SVTK_VCF_NO_INFO_LINES = 1, SVTK_VCF_OWNS_UNPLACED = 2, SVTK_VCF_TEE_STDOUT = 4 /* lines go to out_path AND stdout */,
       SVTK_VCF_QUIET = 8 /* do not repeat the reference's stderr messages */ };
int svtk_process_vcf_region(svtk_ctx *ctx, const char *vcf_path, const char *out_path, int32_t tid_lo, int32_t tid_hi, uint32_t flags,
                            int64_t *n_records);
/* svtk_process_vcf with flags (SVTK_VCF_TEE_STDOUT: the reference accepts -o and never writes it; the C driver treats
 * it as an additional sink) and the optional consensus side file of svtk_process_vcf_ins (fasta_path NULL: none). */
int svtk_process_vcf_flags(svtk_ctx *ctx, const char *vcf_path, const char *out_path, const char *fasta_path, int32_t tol, uint32_t flags,
                           int64_t *n_records, int64_t *n_consensus);
/* One region shard of a whole job, file to file (multi-GPU runs; replaces the worker fan-out of audit.c:269-293): the VCF's records
 * are cut into n_shards contiguous runs (VCF order == genomic order, cuts may fall inside a reference) of equal estimated work --
 * the compressed BAM bytes under each record's windows, read off the .bai's linear index -- and this call handles run `shard`:
 * it loads only what those windows can see (svtk_load_bam_windows), refines them and writes their lines to out_path.  Every
 * shard computes the same cuts; the outputs concatenated in shard order are svtk_process_vcf's output.  One ctx (GPU) per shard. */
int svtk_audit_shard(svtk_ctx *ctx, const char *bam_path, const char *bai_path, const char *vcf_path, const char *out_path, int32_t shard,
                     int32_t n_shards, uint32_t flags, int64_t *n_records, svtk_bam_stats *stats);
/* The front and back ends of process_vcf alone -- no ctx, no GPU: the windows process_vcf builds for a VCF (audit.c:65-193,
 * in file order; call with cap 0 for the count), and the lines it writes for given per-window results (audit.c:182-229).
 * n_ref: size of the BAM's reference dictionary (0: unknown), only used to place records of a sharded call.  tid_lo < 0: the
 * whole file.  Parity aids: the host logic can be checked against the reference's captured stdout without a device. */
int svtk_vcf_windows(const char *vcf_path, const svtk_params *params, int32_t n_ref, int32_t tid_lo, int32_t tid_hi, uint32_t flags,
                     int64_t cap, int32_t *chrom, uint32_t *start, uint32_t *end, uint32_t *pos, uint8_t *kind, int64_t *n_win);
int svtk_vcf_format(const char *vcf_path, const svtk_params *params, int32_t n_ref, int32_t tid_lo, int32_t tid_hi, uint32_t flags,
                    const int32_t *results, int64_t n_results, const char *out_path, int64_t *n_records);
/* The same, plus the side file the north star's "abPOA for insertion consensus" needs (the reference's stdout has no slot for a
 * sequence): for every INS record whose refinement gave a position and that has carrier reads, one FASTA entry
 *   >INS chr=<chrom> org_pos=<VCF pos> ref_pos=<refined> reads=<carriers> len=<bases>
 * holding the POA consensus (svtk_poa_consensus == abpoa_msa defaults) of the inserted sequences svtk_ins_sequences finds
 * within tol bases of the refined position (tol < 0: the ctx's consensus_interval).  Needs reads loaded with svtk_load_bam. */
int svtk_process_vcf_ins(svtk_ctx *ctx, const char *vcf_path, const char *out_path, const char *fasta_path, int32_t tol, int64_t *n_records,
                         int64_t *n_consensus);

/* ---- K3: insertion consensus (partial-order alignment).  SVTrek builds abPOA but never calls it (no
 * source includes abpoa.h, Makefile:32-33), so this entry has no call site in audit.c; it replaces what the
 * north star names "abPOA for insertion consensus": abpoa_msa() with default parameters
 * (abPOA/src/abpoa_align.c:401, abPOA/include/abpoa.h:154) on one read set per INS record.
 * A batch is n_sets independent read sets: set s owns reads set_off[s] .. set_off[s+1], read r owns
 * bases[seq_off[r] .. seq_off[r+1]) with abPOA's nucleotide codes A0 C1 G2 T3 other 4 (abpoa_seq.c:15).
 * Reads are aligned in input order, as abpoa_poa does (abpoa_align.c:312).  Empty reads are an error. */
typedef struct {
    int32_t n_sets;
    const int64_t *set_off;  /* [n_sets + 1] */
    const int64_t *seq_off;  /* [n_reads + 1], seq_off[0] == 0 */
    const uint8_t *bases;    /* [seq_off[n_reads]] */
} svtk_poa_batch;

/* optional per-read observations (host arrays of n_reads entries; any pointer may be NULL only if all are):
 * what abpoa_res_t.best_score, abg->node_n after the read, and the alignment (expanded ops: FNV-1a hash) were. */
typedef struct {
    int32_t *best_score;
    int32_t *node_n;
    int32_t *n_ops;
    uint64_t *ops_hash;
    int64_t *cells;          /* band cells of the read's alignment */
} svtk_poa_trace;

typedef struct {
    uint64_t n_sets, n_reads, n_bases;
    uint64_t cells;          /* 5-state convex cell updates actually computed: sum of band widths over all DP rows */
    uint64_t cells_full;     /* full-matrix equivalent: sum over reads r >= 2 of rows * (qlen + 1) */
    uint64_t relaunched;     /* read sets that outgrew their first scratch estimate and were rerun with a larger one */
    uint64_t workers;        /* resident warps (= CTAs) of the largest launch */
    double kernel_ms;        /* CUDA-event time of the k_poa_sets launches on the ctx stream */
    uint64_t phase_cycles[4];/* SM cycles summed over read sets: row DP, backtrack, graph fusion, topological sort */
} svtk_poa_stats;

enum { SVTK_POA_OK = 0, SVTK_POA_FAILED = 1 };

/* cons / cons_cov: host buffers; set s writes at most cons_off[s+1] - cons_off[s] entries at cons_off[s]
 * (cons_cov may be NULL: abpoa_cons_t.cons_cov, abpoa.h:122).  cons_len[s] receives the full consensus length
 * (abpoa_cons_t.cons_len) even when it exceeds the room; set_status[s] (may be NULL) is SVTK_POA_OK or SVTK_POA_FAILED. */
int svtk_poa_consensus(svtk_ctx *ctx, const svtk_poa_batch *host_batch, uint8_t *cons, const int64_t *cons_off, int32_t *cons_len,
                       int32_t *cons_cov, int32_t *set_status, const svtk_poa_trace *trace, svtk_poa_stats *stats);

#ifdef __cplusplus
}
#endif
#endif /* SVTK_H */
