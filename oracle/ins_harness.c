/* TEST INFRASTRUCTURE ONLY -- the inserted sequences behind a refined insertion breakpoint, read with the REFERENCE'S OWN
 * HTSlib (oracle/_ref/libhts.a, built unmodified from /root/reference/htslib by oracle/build_ref.sh): pins SURVEY 8f-2
 * (svtk_ins_sequences) to the library the reference reads BAMs with, instead of to a Python restatement.
 *
 * Per INS window: the iterator refine_ins opens -- sam_itr_queryi(idx, chrom - 1, start - 1, end - 1) (refinement.c:285) -- and
 * the loop it runs over sam_itr_next (refinement.c:293): every `I` op with len >= min_len (refinement.c:299: 50) met before the
 * reference cursor passes inter.end (refinement.c:314-317), cursor advanced by every op but I and S (refinement.c:310-312).
 * Where refine_ins pushes the cursor, this harness ALSO takes the op's bases with the reference's accessors bam_get_seq /
 * bam_seqi (htslib/sam.h:300,325) when the cursor lies within tol of the window's refined position.  Nothing of refinement.c is
 * reproduced beyond that loop shape; the selection of reads and the decoding of SEQ are HTSlib's. */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <htslib/sam.h>

/* bases come out as abPOA codes: "=ACMGRSVTWYHKDBN"[nibble] -> A0 C1 G2 T3, anything else 4 (abPOA/src/abpoa_seq.c:15) */
static uint8_t code_of(int nib) { return nib == 1 ? 0 : nib == 2 ? 1 : nib == 4 ? 2 : nib == 8 ? 3 : 4; }

/* Returns 0, or -1 if a capacity was too small (n_seqs / n_bases still receive the full counts), -2 on I/O errors.
 * Per sequence: window index, core.pos of the read, reference cursor of the op; seq_off[n_seqs + 1] into bases. */
int ref_ins_sequences(const char *bam_path, int n_win, const int32_t *chrom, const uint32_t *start, const uint32_t *end, const uint8_t *kind,
                      const int32_t *refined, int tol, int min_len, int64_t cap_seqs, int64_t cap_bases, int32_t *win_out, uint32_t *read_pos_out,
                      uint32_t *ref_pos_out, int64_t *seq_off, uint8_t *bases, int64_t *n_seqs, int64_t *n_bases) {
    samFile *fp = sam_open(bam_path, "r");
    if (!fp) return -2;
    bam_hdr_t *hdr = sam_hdr_read(fp);
    hts_idx_t *idx = sam_index_load(fp, bam_path);
    if (!hdr || !idx) { if (idx) hts_idx_destroy(idx); if (hdr) bam_hdr_destroy(hdr); sam_close(fp); return -2; }
    bam1_t *aln = bam_init1();
    int64_t ns = 0, nb = 0;
    int short_of_room = 0;
    for (int w = 0; w < n_win; ++w) {
        if (kind[w] != 2 || refined[w] < 0) continue;
        hts_itr_t *iter = sam_itr_queryi(idx, chrom[w] - 1, start[w] - 1, end[w] - 1);
        if (!iter) continue;
        while (sam_itr_next(fp, iter, aln) > 0) {
            const uint32_t *cigar = bam_get_cigar(aln);
            const uint8_t *seq = bam_get_seq(aln);
            uint32_t reference_pos = (uint32_t)aln->core.pos;
            int64_t q = 0;
            for (uint32_t i = 0; i < aln->core.n_cigar; ++i) {
                const int op = bam_cigar_op(cigar[i]);
                const int len = (int)bam_cigar_oplen(cigar[i]);
                if (op == BAM_CINS && len >= min_len) {
                    int64_t d = (int64_t)reference_pos - (int64_t)refined[w];
                    if (d < 0) d = -d;
                    if (d <= tol && q + len <= aln->core.l_qseq) {
                        if (ns < cap_seqs && nb + len <= cap_bases) {
                            win_out[ns] = w; read_pos_out[ns] = (uint32_t)aln->core.pos; ref_pos_out[ns] = reference_pos; seq_off[ns] = nb;
                            for (int k = 0; k < len; ++k) bases[nb + k] = code_of(bam_seqi(seq, q + k));
                        } else short_of_room = 1;
                        ns += 1; nb += len;
                    }
                }
                if (op != BAM_CINS && op != BAM_CSOFT_CLIP) reference_pos += (uint32_t)len;
                if (bam_cigar_type(op) & 1) q += len;
                if (reference_pos > end[w]) break;
            }
        }
        sam_itr_destroy(iter);
    }
    if (ns <= cap_seqs) seq_off[ns < cap_seqs ? ns : cap_seqs] = nb;
    *n_seqs = ns; *n_bases = nb;
    bam_destroy1(aln);
    hts_idx_destroy(idx);
    bam_hdr_destroy(hdr);
    sam_close(fp);
    return short_of_room ? -1 : 0;
}
