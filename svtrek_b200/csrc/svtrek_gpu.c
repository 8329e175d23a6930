/* svtrek_gpu.c -- `svtrek audt` with the refinement path on the GPU: a host driver in the reference's own language (C) over
 * the C ABI of libsvtk_b200.so.  Same command line as the reference's audt mode (init.c:49-140: -b/--bam, -v/--vcf,
 * -o/--output, -t, --verbose, --wider-interval, --median-interval, --narrow-interval, --consensus-interval-range,
 * --consensus-interval, --consensus-min-count, -h/--help; defaults params.h:27-36), same stdout: the two [INFO] lines of
 * process_vcf (audit.c:252,354) around one (INS)/(DEL)/(INV) line per record (audit.c:182-184,196-217,229), same stderr
 * messages for malformed VCF lines.  Differences, all deliberate: every record is printed, in VCF order (the reference drops
 * up to 2 x t queued lines at end of file and interleaves output at -t > 1: SURVEY F2/F3) -- i.e. the output of
 * `svtrek audt -t 1` on a VCF padded past the tail drop; -t is accepted and ignored (the GPU replaces the thread pool);
 * -o, when given, receives a copy of the record lines in addition to stdout (the reference accepts -o and never writes to
 * it -- SURVEY F1; the additional sink is SURVEY 8f-3's reading of the option).
 * Extra options: --device N (default 0), --devices a,b,c (one host thread + one ctx per listed GPU: the job is cut into region
 * shards of equal work, cuts may fall inside a reference, every GPU loads only the part of the BAM its records' windows can see;
 * shard outputs are concatenated in VCF order -- the north star's replacement for tpool's worker fan-out, audit.c:269-293),
 * --stats (loader counters on stderr), --ins-consensus FILE (FASTA side file with the POA
 * consensus of the inserted sequence of every refined INS record: the north star's "abPOA for insertion consensus").  `disc` mode is not part of this path. */
#include <getopt.h>
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

#include "../../include/svtk.h"

static void usage(void) {
    printf("Usage: ./svtrek_gpu audt [-b|--bam BAM] [-v|--vcf VCF file] [OPTIONS]\n");
    printf("Options:\n");
    printf("    [-o|--ouput] <filename>           Output filename [Default: svtrek.out]\n");
    printf("    -t <num>                          Thread number (accepted, unused: the GPU path has no thread pool)\n");
    printf("    --verbose                         Verbose [Default: false]\n\n");
    printf("    --wider-interval <num>            Interval for the offset of the reads to start [Default: 20000]\n");
    printf("    --median-interval <num>           Interval for the offset of the reads (for point) [Default: 10000]\n");
    printf("    --narrow-interval <num>           Interval for the offset of the reads to end [Default: 2000]\n");
    printf("    --consensus-interval-range <num>  The interval to limit refinement range [DEFAULT: 500]\n");
    printf("    --consensus-interval <num>        The interval that is considered into the same position [DEFAULT: 5]\n");
    printf("    --consensus-min-count <num>       Minimum number of elements needs for the consensus [Default: 3]\n\n");
    printf("    --device <num>                    CUDA device [Default: 0]\n");
    printf("    --devices <a,b,...>               Several CUDA devices: the job is cut into one region shard per device\n");
    printf("    --stats                           BAM loader counters on stderr\n");
    printf("    --ins-consensus <filename>        FASTA with the consensus of the inserted sequence of every refined INS record\n\n");
}

static void need_file(const char *name, const char *missing) {
    FILE *f;
    if (!name) { fprintf(stderr, "%s\n", missing); exit(EXIT_FAILURE); }
    f = fopen(name, "r");
    if (!f) { fprintf(stderr, "[ERROR]: File couldn't be opened %s\n", name); exit(EXIT_FAILURE); }
    fclose(f);
}

/* one region shard on its own GPU, from its own host thread (one ctx per thread, like the reference's one t_arg per worker) */
typedef struct {
    int device, shard, n_shards, stats, rc;
    const char *bam, *vcf;
    svtk_params p;
    char out[64];
    long long n;
    char err[512];
} shard_job;

static void *shard_main(void *arg) {
    shard_job *j = (shard_job *)arg;
    svtk_bam_stats st;
    int64_t n = 0;
    svtk_ctx *ctx = svtk_create(j->device, &j->p, NULL);
    j->rc = -1;
    if (!ctx) { snprintf(j->err, sizeof j->err, "%s", svtk_last_error()); return NULL; }
    j->rc = svtk_audit_shard(ctx, j->bam, NULL, j->vcf, j->out, j->shard, j->n_shards, SVTK_VCF_NO_INFO_LINES, &n, &st);
    if (j->rc) snprintf(j->err, sizeof j->err, "%s", svtk_last_error());
    else if (j->stats)
        fprintf(stderr, "[svtrek_gpu] shard %d/%d on device %d: shipped %.1f of %.1f MB in %u span(s), %llu BGZF blocks, %llu reads; %lld records; upload %.1f ms, "
                        "inflate %.1f ms, total %.1f ms\n", j->shard, j->n_shards, j->device, st.image_bytes / 1e6, st.file_bytes / 1e6, st.n_spans,
                (unsigned long long)st.n_blocks, (unsigned long long)st.n_reads, (long long)n, st.upload_ms, st.inflate_ms, st.total_ms);
    j->n = (long long)n;
    svtk_destroy(ctx);
    return NULL;
}

int main(int argc, char *argv[]) {
    const char *bam = NULL, *vcf = NULL, *out = NULL, *fasta = NULL, *devlist = NULL;
    svtk_params p = {20000, 10000, 2000, 500, 5, 3};
    int device = 0, stats = 0, verbose = 0, opt, li, threads = 4;
    static struct option lo[] = {
        {"bam", required_argument, NULL, 1}, {"vcf", required_argument, NULL, 2}, {"output", required_argument, NULL, 3},
        {"verbose", no_argument, NULL, 4}, {"wider-interval", required_argument, NULL, 5}, {"median-interval", required_argument, NULL, 6},
        {"narrow-interval", required_argument, NULL, 7}, {"consensus-interval-range", required_argument, NULL, 8},
        {"consensus-interval", required_argument, NULL, 9}, {"consensus-min-count", required_argument, NULL, 10},
        {"help", no_argument, NULL, 11}, {"device", required_argument, NULL, 12}, {"stats", no_argument, NULL, 13}, {"ins-consensus", required_argument, NULL, 14},
        {"devices", required_argument, NULL, 15}, {NULL, 0, NULL, 0}};
    if (argc < 2 || strcmp(argv[1], "audt") != 0) {
        printf("Usage: ./svtrek_gpu audt [OPTIONS]   (only the audt mode runs on the GPU)\n");
        return 1;
    }
    if (argc < 3) { usage(); return 1; }
    argc -= 1; argv += 1;                                   /* like audit(): getopt sees "audt" as argv[0] */
    while ((opt = getopt_long(argc, argv, "b:v:o:t:h", lo, &li)) != -1) {
        switch (opt) {
            case 'b': case 1: bam = optarg; break;
            case 'v': case 2: vcf = optarg; break;
            case 'o': case 3: out = optarg; break;
            case 4: verbose = 1; break;
            case 't': threads = atoi(optarg); break;
            case 5: p.wider_interval = atoi(optarg); break;
            case 6: p.median_interval = atoi(optarg); break;
            case 7: p.narrow_interval = atoi(optarg); break;
            case 8: p.consensus_interval_range = atoi(optarg); break;
            case 9: p.consensus_interval = atoi(optarg); break;
            case 10: p.consensus_min_count = atoi(optarg); break;
            case 'h': case 11: usage(); return EXIT_SUCCESS;
            case 12: device = atoi(optarg); break;
            case 13: stats = 1; break;
            case 14: fasta = optarg; break;
            case 15: devlist = optarg; break;
            default:
                printf("[ERROR] Option %d is invalid.\n", opt);
                usage();
                return EXIT_FAILURE;
        }
    }
    (void)verbose; (void)threads;
    need_file(bam, "[ERROR] BAM file is not provided.");
    need_file(vcf, "[ERROR] VCF file is not provided.");

    if (devlist) {                                          /* ---- several GPUs: one region shard each, merged in VCF order */
        shard_job jobs[64];
        pthread_t th[64];
        int nd = 0, k, bad = 0;
        long long total = 0;
        char *list = strdup(devlist), *save = NULL, *tok;
        FILE *fo = out ? fopen(out, "w") : NULL;
        if (fasta) fprintf(stderr, "[svtrek_gpu] --ins-consensus runs on one device: ignored with --devices\n");
        for (tok = strtok_r(list, ",", &save); tok && nd < 64; tok = strtok_r(NULL, ",", &save)) jobs[nd++].device = atoi(tok);
        if (nd == 0) { fprintf(stderr, "[ERROR] --devices needs a list\n"); return EXIT_FAILURE; }
        for (k = 0; k < nd; ++k) {
            jobs[k].shard = k; jobs[k].n_shards = nd; jobs[k].stats = stats; jobs[k].bam = bam; jobs[k].vcf = vcf; jobs[k].p = p; jobs[k].n = 0; jobs[k].err[0] = 0;
            snprintf(jobs[k].out, sizeof jobs[k].out, "/tmp/svtrek_gpu.%d.%d.out", (int)getpid(), k);
            pthread_create(&th[k], NULL, shard_main, &jobs[k]);
        }
        printf("[INFO] Started processing variation file.\n");      /* audit.c:252 */
        if (fo) fputs("[INFO] Started processing variation file.\n", fo);
        for (k = 0; k < nd; ++k) {
            pthread_join(th[k], NULL);
            if (jobs[k].rc != 0) { fprintf(stderr, "[ERROR] shard %d (device %d): %s\n", k, jobs[k].device, jobs[k].err); bad = 1; }
            else {
                FILE *f = fopen(jobs[k].out, "r");
                char buf[1 << 16];
                size_t got;
                if (f) { while ((got = fread(buf, 1, sizeof buf, f)) > 0) { fwrite(buf, 1, got, stdout); if (fo) fwrite(buf, 1, got, fo); } fclose(f); }
                total += jobs[k].n;
            }
            unlink(jobs[k].out);
        }
        if (!bad) { printf("[INFO] Ended processing variation file\n"); if (fo) fputs("[INFO] Ended processing variation file\n", fo); }   /* audit.c:354 */
        if (fo) fclose(fo);
        if (stats) fprintf(stderr, "[svtrek_gpu] %lld records on %d devices\n", total, nd);
        free(list);
        return 0;
    }

    svtk_ctx *ctx = svtk_create(device, &p, NULL);
    if (!ctx) { fprintf(stderr, "[ERROR] %s\n", svtk_last_error()); return EXIT_FAILURE; }
    svtk_bam_stats st;
    if (svtk_load_bam(ctx, bam, NULL, &st) != 0) {
        fprintf(stderr, "[ERROR] %s\n", svtk_last_error());
        svtk_destroy(ctx);
        return EXIT_FAILURE;
    }
    if (stats)
        fprintf(stderr, "[svtrek_gpu] %s: %.1f MB, %llu BGZF blocks -> %.1f MB, %llu reads, %llu CIGAR words; upload %.1f ms, inflate %.1f ms (both: %.1f ms), "
                        "records %.2f ms, total %.1f ms\n", bam, st.file_bytes / 1e6, (unsigned long long)st.n_blocks, st.inflated_bytes / 1e6,
                (unsigned long long)st.n_reads, (unsigned long long)st.n_words, st.upload_ms, st.inflate_ms, st.load_span_ms, st.parse_ms, st.total_ms);
    int64_t n = 0, n_cons = 0;
    /* -o is an additional sink: the lines go to the file AND to stdout, written once (SVTK_VCF_TEE_STDOUT) */
    int rc = svtk_process_vcf_flags(ctx, vcf, out, fasta, -1, out ? SVTK_VCF_TEE_STDOUT : 0, &n, &n_cons);
    if (rc != 0) fprintf(stderr, "[ERROR] %s\n", svtk_last_error());
    if (stats) fprintf(stderr, "[svtrek_gpu] %lld records, %lld insertion consensus sequences\n", (long long)n, (long long)n_cons);
    svtk_destroy(ctx);
    return 0;                                               /* the reference's exit status is 0 whatever happened (svtrek.c:15-21) */
}
