// svtk_host.cpp -- host side of the batch seam: the VCF front end and the stdout back end of process_vcf
// (audit.c:250-357) around the GPU refinement.  thread_func's per-line parsing (audit.c:65-173) and its printf
// formats (audit.c:182-229) are reproduced byte for byte; the tpool fan-out (audit.c:287-293) is replaced by
// one batched launch per chunk of records, results are written in VCF order.
#include "../../include/svtk.h"
#include "svtk_ctx.h"

#include <algorithm>
#include <cerrno>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

namespace svtk {

enum SvType { T_UNKNOWN = 0, T_INS, T_DEL, T_INV, T_DUP, T_TRA, T_BND };   // params.h:113-121

struct Record { int type; int chrom; uint32_t pos, end; };

// strtok_r-style splitter: tokens are maximal runs of non-delimiter bytes, terminated in place
struct Splitter {
    char *cur;
    char delim;
    char *next() {
        while (*cur == delim) ++cur;
        if (!*cur) return nullptr;
        char *tok = cur;
        while (*cur && *cur != delim) ++cur;
        if (*cur) *cur++ = 0;
        return tok;
    }
};

static int type_of(const char *s) {                        // audit.c:3-11
    struct { const char *name; int t; } tab[] = {{"INS", T_INS}, {"INS:ME", T_INS}, {"DEL", T_DEL}, {"DEL:ME", T_DEL},
                                                 {"INV", T_INV}, {"DUP", T_DUP}, {"TRA", T_TRA}, {"BND", T_BND}};
    for (auto &e : tab) if (!strcmp(s, e.name)) return e.t;
    return T_UNKNOWN;
}

// value of KEY in an INFO string the way the reference extracts it: first strstr hit, up to ';', clipped to cap-1
static bool info_value(const char *info, const char *key, char *out, size_t cap) {
    const char *p = strstr(info, key);
    if (!p) return false;
    p += strlen(key);
    const char *e = strchr(p, ';');
    size_t len = e ? (size_t)(e - p) : strlen(p);
    if (len >= cap) len = cap - 1;
    memcpy(out, p, len);
    out[len] = 0;
    return true;
}

// 1: record to audit, 0: skipped, -1: "[ERROR] Unkown type."
static int parse_record(char *line, Record *r) {
    Splitter f{line, '\t'};
    char *chrom = f.next();
    char *index = f.next();
    if (!index) { fprintf(stderr, "VCF: no index at line: %s\n", chrom ? chrom : line); return 0; }   // audit.c:67-71
    f.next();                                              // ID
    char *ref = f.next();
    char *alt = f.next();
    if (!ref || !alt) return 0;                            // (the reference dereferences NULL here)
    const size_t ref_len = strlen(ref);
    size_t max_alt = 0, min_alt = 0x7FFFFFFF;              // audit.c:80-90
    Splitter a{alt, ','};
    for (char *t = a.next(); t; t = a.next()) {
        const size_t l = strlen(t);
        if (l > max_alt) max_alt = l;
        if (l < min_alt) min_alt = l;
    }
    f.next(); f.next();                                    // QUAL, FILTER
    char *info = f.next();
    if (!info) return 0;

    r->chrom = !strncmp(chrom, "chr", 3) ? atoi(chrom + 3) : atoi(chrom);      // audit.c:101-105
    r->pos = (uint32_t)strtol(index, nullptr, 10);                             // audit.c:108
    if (r->pos == 0 && index[0] != '0') { fprintf(stderr, "[ERROR] Conversion error to pos %s\n", index); return 0; }

    char tbuf[16], ebuf[32];
    if (info_value(info, "SVTYPE=", tbuf, sizeof tbuf)) r->type = type_of(tbuf);               // audit.c:116-131
    else if ((int)ref_len == 1 && 50 < max_alt) r->type = T_INS;                               // audit.c:133-141
    else if (50 < (int)ref_len && min_alt == 1) r->type = T_DEL;
    else return 0;

    if (info_value(info, "END=", ebuf, sizeof ebuf)) {     // first hit -- may be the tail of "CIEND=" (audit.c:145)
        r->end = (uint32_t)strtol(ebuf, nullptr, 10);
        if (r->end == 0 && ebuf[0] != '0') return 0;
    } else r->end = r->pos + (uint32_t)ref_len;            // audit.c:165

    if ((r->type == T_DEL || r->type == T_INV) && r->end - r->pos < 50u) return 0;             // audit.c:168-173
    if (r->type == T_INS || r->type == T_DEL || r->type == T_INV) return 1;
    fprintf(stderr, "[ERROR] Unkown type.\n");             // audit.c:233-235
    return -1;
}

struct Batch {
    std::vector<Record> rec;
    std::vector<int32_t> chrom;
    std::vector<uint32_t> start, end, pos;
    std::vector<uint8_t> kind;
    std::vector<int32_t> first;                            // index of each record's first window (-1: prints nothing)
    void clear() { rec.clear(); chrom.clear(); start.clear(); end.clear(); pos.clear(); kind.clear(); first.clear(); }
    void window(int c, uint32_t s, uint32_t e, uint32_t p, int k) {
        chrom.push_back(c); start.push_back(s); end.push_back(e); pos.push_back(p); kind.push_back((uint8_t)k);
    }
};

// window construction: audit.c:178 (INS), 191-193 (DEL), 224-226 (INV); uint32 arithmetic wraps like the reference's
static void add_record(Batch &b, const Record &r, const svtk_params &p) {
    int first = -1;
    if (r.type == T_INS) {
        first = (int)b.chrom.size();
        b.window(r.chrom, r.pos - (uint32_t)p.median_interval, r.pos + (uint32_t)p.median_interval, r.pos, SVTK_INS);
    } else if (50u < r.end - r.pos) {
        first = (int)b.chrom.size();
        if (r.type == T_DEL) {
            b.window(r.chrom, r.pos - (uint32_t)p.wider_interval, r.pos + (uint32_t)p.narrow_interval, r.pos, SVTK_DEL_START);
            b.window(r.chrom, r.end - (uint32_t)p.narrow_interval, r.end + (uint32_t)p.narrow_interval, r.end, SVTK_DEL_END);
        } else {
            b.window(r.chrom, r.pos - (uint32_t)p.wider_interval, r.pos + (uint32_t)p.wider_interval, r.pos, SVTK_POINT);
            b.window(r.chrom, r.end - (uint32_t)p.wider_interval, r.end + (uint32_t)p.wider_interval, r.end, SVTK_POINT);
        }
    }
    b.rec.push_back(r);
    b.first.push_back(first);
}

static void format_record(std::string &out, const Record &r, const int32_t *res) {
    char buf[256];
    const uint32_t NA = 0xFFFFFFFFu;
    if (r.type == T_INS) {                                 // audit.c:181-185
        const uint32_t v = (uint32_t)res[0];
        if (v == NA) snprintf(buf, sizeof buf, "(INS) chr: %d, org pos: %u, ref pos: NA\n", r.chrom, r.pos);
        else snprintf(buf, sizeof buf, "(INS) chr: %d, org pos: %u, ref pos: %u, diff: %d\n", r.chrom, r.pos, v, (int)(v - r.pos));
        out += buf;
    } else if (r.type == T_DEL) {                          // audit.c:196-217
        const uint32_t s = (uint32_t)res[0], e = (uint32_t)res[1];
        snprintf(buf, sizeof buf, "(DEL) chr: %d, org pos: %u, org end: %u, ref pos: ", r.chrom, r.pos, r.end);
        out += buf;
        if (s == NA) out += "NA, ref end: "; else { snprintf(buf, sizeof buf, "%d, ref end: ", (int)s); out += buf; }
        if (e == NA) out += "NA, "; else { snprintf(buf, sizeof buf, "%d, ", (int)e); out += buf; }
        if (s == NA) out += "diff pos: NA, "; else { snprintf(buf, sizeof buf, "diff pos: %d, ", (int)(s - r.pos)); out += buf; }
        if (e == NA) out += "diff end: NA\n"; else { snprintf(buf, sizeof buf, "diff end: %d\n", (int)(e - r.end)); out += buf; }
    } else {                                               // audit.c:229
        snprintf(buf, sizeof buf, "(INV) chr: %d, org pos: %u, org end: %u, ref pos: %u, ref end: %u\n", r.chrom, r.pos, r.end,
                 (uint32_t)res[0], (uint32_t)res[1]);
        out += buf;
    }
}

// the inserted sequences of the batch's refined INS records -> POA consensus -> FASTA (one entry per record that has carriers)
static int write_ins_consensus(svtk_ctx *ctx, const Batch &b, const svtk_windows &w, const int32_t *res, int tol, FILE *fa, int64_t *n_cons) {
    svtk_ins_sets sets;
    if (int rc = svtk_ins_sequences(ctx, &w, res, tol, 50, &sets)) return rc;
    if (sets.n_sets == 0) return 0;
    std::vector<int> rec_of_win(b.chrom.size(), -1);
    for (size_t i = 0; i < b.rec.size(); ++i) if (b.first[i] >= 0) rec_of_win[(size_t)b.first[i]] = (int)i;
    std::vector<int64_t> cons_off((size_t)sets.n_sets + 1, 0);
    for (int s = 0; s < sets.n_sets; ++s) {
        int64_t longest = 0;
        for (int64_t k = sets.set_off[s]; k < sets.set_off[s + 1]; ++k) longest = std::max(longest, sets.seq_off[k + 1] - sets.seq_off[k]);
        cons_off[(size_t)s + 1] = cons_off[(size_t)s] + 2 * longest + 64;
    }
    std::vector<uint8_t> cons((size_t)cons_off.back());
    std::vector<int32_t> clen((size_t)sets.n_sets), status((size_t)sets.n_sets);
    // the set arrays belong to the ctx and svtk_poa_consensus does not touch them
    const svtk_poa_batch pb = {sets.n_sets, sets.set_off, sets.seq_off, sets.bases};
    if (int rc = svtk_poa_consensus(ctx, &pb, cons.data(), cons_off.data(), clen.data(), nullptr, status.data(), nullptr, nullptr)) return rc;
    std::string text;
    for (int s = 0; s < sets.n_sets; ++s) {
        const int win = sets.window[s], ri = rec_of_win[(size_t)win];
        if (ri < 0 || status[(size_t)s] != 0) continue;
        const int64_t room = cons_off[(size_t)s + 1] - cons_off[(size_t)s];
        const int32_t n = (int32_t)std::min<int64_t>(clen[(size_t)s], room);
        char head[160];
        snprintf(head, sizeof head, ">INS chr=%d org_pos=%u ref_pos=%d reads=%lld len=%d\n", b.rec[(size_t)ri].chrom, b.rec[(size_t)ri].pos, res[win],
                 (long long)(sets.set_off[s + 1] - sets.set_off[s]), n);
        text += head;
        const uint8_t *q = cons.data() + cons_off[(size_t)s];
        for (int32_t k = 0; k < n; ++k) text += "ACGTN"[q[k] < 4 ? q[k] : 4];
        text += '\n';
        ++*n_cons;
    }
    fwrite(text.data(), 1, text.size(), fa);
    return 0;
}

static int flush_batch(svtk_ctx *ctx, Batch &b, FILE *out, int64_t *n_lines, FILE *fa = nullptr, int tol = 0, int64_t *n_cons = nullptr) {
    if (b.rec.empty()) return 0;
    std::vector<int32_t> res(b.chrom.size());
    const svtk_windows w = {(int32_t)b.chrom.size(), b.chrom.data(), b.start.data(), b.end.data(), b.pos.data(), b.kind.data()};
    if (int rc = svtk_refine_windows(ctx, &w, res.data())) return rc;
    if (fa) if (int rc = write_ins_consensus(ctx, b, w, res.data(), tol, fa, n_cons)) return rc;
    std::string text;
    text.reserve(b.rec.size() * 96);
    for (size_t i = 0; i < b.rec.size(); ++i) {
        if (b.first[i] < 0) continue;
        format_record(text, b.rec[i], res.data() + b.first[i]);
        ++*n_lines;
    }
    fwrite(text.data(), 1, text.size(), out);
    b.clear();
    return 0;
}

}  // namespace svtk

using namespace svtk;

static int process_vcf_impl(svtk_ctx *ctx, const char *vcf_path, const char *out_path, const char *fasta_path, int tol, int64_t *n_records,
                            int64_t *n_consensus) {
    Ctx *c = reinterpret_cast<Ctx *>(ctx);
    if (!c || !vcf_path) return fail(SVTK_ERR_ARG, "null argument");
    if (fasta_path && !c->reads_from_bam) return fail(SVTK_ERR_STATE, "insertion consensus needs the BAM image: load the reads with svtk_load_bam");
    FILE *fa = fasta_path ? fopen(fasta_path, "w") : nullptr;
    if (fasta_path && !fa) return fail(SVTK_ERR_IO, "cannot open %s: %s", fasta_path, strerror(errno));
    struct Closer { FILE *f; ~Closer() { if (f) fclose(f); } } fa_closer{fa};
    int64_t n_cons = 0;
    if (!c->have_reads) return fail(SVTK_ERR_STATE, "no reads loaded (svtk_load_reads)");
    FILE *vf = fopen(vcf_path, "r");
    if (!vf) return fail(SVTK_ERR_IO, "cannot open %s: %s", vcf_path, strerror(errno));
    const bool to_stdout = !out_path || !strcmp(out_path, "-");
    FILE *out = to_stdout ? stdout : fopen(out_path, "w");
    if (!out) { fclose(vf); return fail(SVTK_ERR_IO, "cannot open %s: %s", out_path, strerror(errno)); }
    fputs("[INFO] Started processing variation file.\n", out);                 // audit.c:252
    Batch b;
    int64_t n_lines = 0;
    char *line = nullptr;
    size_t cap = 0;
    ssize_t len;
    int rc = 0;
    const size_t BATCH_RECORDS = 1u << 20;
    while ((len = getline(&line, &cap, vf)) >= 0) {
        if (len < 2 || line[0] == '#') continue;                                // audit.c:324
        if (line[len - 1] == '\n') line[--len] = 0;                            // audit.c:327-330
        Record r;
        if (parse_record(line, &r) == 1) add_record(b, r, c->params);
        if (b.rec.size() >= BATCH_RECORDS && (rc = flush_batch(ctx, b, out, &n_lines, fa, tol, &n_cons))) break;
    }
    if (!rc) rc = flush_batch(ctx, b, out, &n_lines, fa, tol, &n_cons);
    free(line);
    fclose(vf);
    if (!rc) fputs("[INFO] Ended processing variation file\n", out);           // audit.c:354
    if (to_stdout) fflush(out); else fclose(out);
    if (n_records) *n_records = n_lines;
    if (n_consensus) *n_consensus = n_cons;
    return rc;
}

extern "C" int svtk_process_vcf(svtk_ctx *ctx, const char *vcf_path, const char *out_path, int64_t *n_records) {
    return process_vcf_impl(ctx, vcf_path, out_path, nullptr, 0, n_records, nullptr);
}

extern "C" int svtk_process_vcf_ins(svtk_ctx *ctx, const char *vcf_path, const char *out_path, const char *fasta_path, int32_t tol, int64_t *n_records,
                                    int64_t *n_consensus) {
    if (!fasta_path) return fail(SVTK_ERR_ARG, "null fasta path");
    if (tol < 0) tol = reinterpret_cast<Ctx *>(ctx) ? reinterpret_cast<Ctx *>(ctx)->params.consensus_interval : 0;
    return process_vcf_impl(ctx, vcf_path, out_path, fasta_path, tol, n_records, n_consensus);
}
