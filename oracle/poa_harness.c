/* TEST INFRASTRUCTURE ONLY -- direct-call shim over the UNMODIFIED abPOA library (reference submodule
 * abPOA/ v1.5.4, compiled from /root/reference/abPOA/src by oracle/build_ref.sh into
 * oracle/_ref/libabpoa_ref.so).  SVTrek itself never calls abPOA, so the insertion-consensus
 * kernel (K3) is pinned at the abPOA library boundary: same byte sequences in, default abpoa_init_para()
 * + abpoa_post_set_para(), and
 *   ref_poa_msa    = abpoa_msa()                                            abpoa_align.c:401
 *   ref_poa_steps  = the same default path unrolled read by read            abpoa_align.c:312-352
 *                    (abpoa_align_sequence_to_graph + abpoa_add_graph_alignment, pattern of abPOA/sub_example.c)
 *                    so the tests can see every read's best_score, alignment and band.
 * No abPOA source is reproduced here. */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <stdint.h>
#include "abpoa.h"

void abpoa_generate_consensus(abpoa_t *ab, abpoa_para_t *abpt);   /* abpoa_output.c:1300 */

static abpoa_para_t *mk_para(void) {
    abpoa_para_t *abpt = abpoa_init_para();
    abpoa_post_set_para(abpt);
    return abpt;
}

static int copy_cons(abpoa_t *ab, uint8_t *cons, int cons_cap, int *cons_cov) {
    abpoa_cons_t *abc = ab->abc;
    if (abc->n_cons < 1) return 0;
    int n = abc->cons_len[0], i;
    for (i = 0; i < n && i < cons_cap; ++i) {
        cons[i] = abc->cons_base[0][i];
        if (cons_cov) cons_cov[i] = abc->cons_cov[0][i];
    }
    return n;
}

/* abpoa_msa on n sequences (codes 0..4, concatenated, seq i = bases[off[i] .. off[i+1])).  Returns the consensus length. */
int ref_poa_msa(int n, const int64_t *off, const uint8_t *bases, uint8_t *cons, int cons_cap, int *cons_cov) {
    abpoa_t *ab = abpoa_init();
    abpoa_para_t *abpt = mk_para();
    int *lens = (int *)malloc(sizeof(int) * (size_t)n);
    uint8_t **seqs = (uint8_t **)malloc(sizeof(uint8_t *) * (size_t)n);
    int i;
    for (i = 0; i < n; ++i) { lens[i] = (int)(off[i + 1] - off[i]); seqs[i] = (uint8_t *)(bases + off[i]); }
    abpoa_msa(ab, abpt, n, NULL, lens, seqs, NULL, NULL);
    int len = copy_cons(ab, cons, cons_cap, cons_cov);
    free(lens); free(seqs);
    abpoa_free(ab); abpoa_free_para(abpt);
    return len;
}

/* 64-bit FNV-1a over the expanded alignment ops: (op, node id) for M / D, (op, -1) per inserted base */
static uint64_t fnv(uint64_t h, int32_t v) {
    int k;
    for (k = 0; k < 4; ++k) { h ^= (uint64_t)((uint32_t)v >> (8 * k) & 0xffu); h *= 1099511628211ull; }
    return h;
}

/* Same path read by read.  Per read r: best_score[r], node_n[r] (graph size after the read was added),
 * n_ops[r] (expanded op count), ops_hash[r], cells[r] = sum over DP rows of (dp_end - dp_beg + 1) as abPOA
 * left them in abm->dp_beg/dp_end.  dump_read >= 0: that read's expanded ops go to dump (pairs op,node).  */
int ref_poa_steps(int n, const int64_t *off, const uint8_t *bases, int32_t *best_score, int32_t *node_n, int32_t *n_ops,
                  uint64_t *ops_hash, int64_t *cells, uint8_t *cons, int cons_cap, int *cons_cov, int dump_read, int32_t *dump,
                  int dump_cap) {
    abpoa_t *ab = abpoa_init();
    abpoa_para_t *abpt = mk_para();
    int i, k;
    abpoa_reset(ab, abpt, 1024);
    ab->abs->n_seq = n;
    for (i = 0; i < n; ++i) {
        int qlen = (int)(off[i + 1] - off[i]);
        uint8_t *q = (uint8_t *)(bases + off[i]);
        abpoa_res_t res;
        res.graph_cigar = 0; res.n_cigar = 0; res.best_score = 0;
        int rows = ab->abg->node_n;
        int aligned = abpoa_align_sequence_to_graph(ab, abpt, q, qlen, &res) >= 0;
        uint64_t h = 1469598103934665603ull;
        int nops = 0;
        int64_t c = 0;
        if (aligned) {
            for (k = 0; k < rows - 1; ++k) c += (int64_t)(ab->abm->dp_end[k] - ab->abm->dp_beg[k] + 1);
            for (k = 0; k < res.n_cigar; ++k) {
                abpoa_cigar_t cg = res.graph_cigar[k];
                int op = (int)(cg & 0xf), l;
                if (op == ABPOA_CMATCH || op == ABPOA_CDEL) {
                    int node = (int)((cg >> 34) & 0x3fffffff);
                    int len = op == ABPOA_CDEL ? (int)((cg >> 4) & 0x3fffffff) : 1;
                    for (l = 0; l < len; ++l) {
                        h = fnv(fnv(h, op), node);
                        if (i == dump_read && 2 * nops + 1 < dump_cap) { dump[2 * nops] = op; dump[2 * nops + 1] = node; }
                        ++nops;
                    }
                } else {
                    int len = (int)((cg >> 4) & 0x3fffffff);
                    for (l = 0; l < len; ++l) {
                        h = fnv(fnv(h, op), -1);
                        if (i == dump_read && 2 * nops + 1 < dump_cap) { dump[2 * nops] = op; dump[2 * nops + 1] = -1; }
                        ++nops;
                    }
                }
            }
        }
        best_score[i] = aligned ? res.best_score : 0;
        n_ops[i] = nops; ops_hash[i] = h; cells[i] = c;
        abpoa_add_graph_alignment(ab, abpt, q, NULL, qlen, NULL, res, i, n, 1);
        node_n[i] = ab->abg->node_n;
        if (res.n_cigar) free(res.graph_cigar);
    }
    abpoa_generate_consensus(ab, abpt);
    int len = copy_cons(ab, cons, cons_cap, cons_cov);
    abpoa_free(ab); abpoa_free_para(abpt);
    return len;
}
