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
 * Extra options: --device N (default 0), --stats (loader counters on stderr), --ins-consensus FILE (FASTA side file with the POA
 * consensus of the inserted sequence of every refined INS record: the north star's "abPOA for insertion consensus").  `disc` mode is not part of this path. */
#include <getopt.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

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

int main(int argc, char *argv[]) {
    const char *bam = NULL, *vcf = NULL, *out = NULL, *fasta = NULL;
    svtk_params p = {20000, 10000, 2000, 500, 5, 3};
    int device = 0, stats = 0, verbose = 0, opt, li, threads = 4;
    static struct option lo[] = {
        {"bam", required_argument, NULL, 1}, {"vcf", required_argument, NULL, 2}, {"output", required_argument, NULL, 3},
        {"verbose", no_argument, NULL, 4}, {"wider-interval", required_argument, NULL, 5}, {"median-interval", required_argument, NULL, 6},
        {"narrow-interval", required_argument, NULL, 7}, {"consensus-interval-range", required_argument, NULL, 8},
        {"consensus-interval", required_argument, NULL, 9}, {"consensus-min-count", required_argument, NULL, 10},
        {"help", no_argument, NULL, 11}, {"device", required_argument, NULL, 12}, {"stats", no_argument, NULL, 13}, {"ins-consensus", required_argument, NULL, 14},
        {NULL, 0, NULL, 0}};
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
            default:
                printf("[ERROR] Option %d is invalid.\n", opt);
                usage();
                return EXIT_FAILURE;
        }
    }
    (void)verbose; (void)threads;
    need_file(bam, "[ERROR] BAM file is not provided.");
    need_file(vcf, "[ERROR] VCF file is not provided.");

    svtk_ctx *ctx = svtk_create(device, &p, NULL);
    if (!ctx) { fprintf(stderr, "[ERROR] %s\n", svtk_last_error()); return EXIT_FAILURE; }
    svtk_bam_stats st;
    if (svtk_load_bam(ctx, bam, NULL, &st) != 0) {
        fprintf(stderr, "[ERROR] %s\n", svtk_last_error());
        svtk_destroy(ctx);
        return EXIT_FAILURE;
    }
    if (stats)
        fprintf(stderr, "[svtrek_gpu] %s: %.1f MB, %llu BGZF blocks -> %.1f MB, %llu reads, %llu CIGAR words; upload %.1f ms, inflate %.1f ms, "
                        "records %.2f ms, total %.1f ms\n", bam, st.file_bytes / 1e6, (unsigned long long)st.n_blocks, st.inflated_bytes / 1e6,
                (unsigned long long)st.n_reads, (unsigned long long)st.n_words, st.upload_ms, st.inflate_ms, st.parse_ms, st.total_ms);
    int64_t n = 0, n_cons = 0;
    int rc = fasta ? svtk_process_vcf_ins(ctx, vcf, out, fasta, -1, &n, &n_cons)
                   : svtk_process_vcf(ctx, vcf, out, &n);    /* out == NULL: stdout */
    if (rc != 0) fprintf(stderr, "[ERROR] %s\n", svtk_last_error());
    if (out && rc == 0) {                                    /* -o is an additional sink: the lines still go to stdout */
        FILE *f = fopen(out, "r");
        char buf[1 << 16];
        size_t k;
        if (f) { while ((k = fread(buf, 1, sizeof buf, f)) > 0) fwrite(buf, 1, k, stdout); fclose(f); }
    }
    if (stats) fprintf(stderr, "[svtrek_gpu] %lld records, %lld insertion consensus sequences\n", (long long)n, (long long)n_cons);
    svtk_destroy(ctx);
    return 0;                                               /* the reference's exit status is 0 whatever happened (svtrek.c:15-21) */
}
