// svtk_host.cpp -- host side of the batch seam: the VCF front end and the stdout back end of process_vcf
// (audit.c:250-357) around the GPU refinement.  thread_func's per-line parsing (audit.c:65-173) and its printf
// formats (audit.c:182-229) are reproduced byte for byte; the tpool fan-out (audit.c:287-293) is replaced by
// one batched launch per chunk of records, results are written in VCF order.
#include "../../include/svtk.h"
#include "svtk_ctx.h"

#include <algorithm>
#include <cerrno>
#include <thread>

#include <fcntl.h>
#include <sys/stat.h>
#include <unistd.h>
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

// 1: record to audit, 0: skipped, -1: "[ERROR] Unkown type."  A message the reference writes to stderr for this line is left
// in `msg` (the caller prints it: lines are parsed by several threads, messages come out in file order)
static int parse_record(char *line, Record *r, std::string *msg) {
    Splitter f{line, '\t'};
    char *chrom = f.next();
    char *index = f.next();
    if (!index) { *msg = std::string("VCF: no index at line: ") + (chrom ? chrom : line) + "\n"; return 0; }   // audit.c:67-71
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
    if (r->pos == 0 && index[0] != '0') { *msg = std::string("[ERROR] Conversion error to pos ") + index + "\n"; return 0; }

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
    *msg = "[ERROR] Unkown type.\n";                       // audit.c:233-235
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
    std::vector<int32_t> clen((size_t)sets.n_sets), status((size_t)sets.n_sets, SVTK_POA_FAILED);
    // the set arrays belong to the ctx and svtk_poa_consensus does not touch them
    const svtk_poa_batch pb = {sets.n_sets, sets.set_off, sets.seq_off, sets.bases};
    if (int rc = svtk_poa_consensus(ctx, &pb, cons.data(), cons_off.data(), clen.data(), nullptr, status.data(), nullptr, nullptr)) {
        // a read set that fails its alignment loses its FASTA entry, nothing else (the breakpoint lines do not depend on K3)
        if (rc != SVTK_ERR_STATE) return rc;
        fprintf(stderr, "[WARN] %s: no consensus for those records\n", svtk_last_error());
    }
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

// one (INS)/(DEL)/(INV) line per record that prints, formatted by several threads, written in VCF order
static void format_all(const Batch &b, const int32_t *res, int threads, std::vector<std::string> &parts, int64_t *n_lines) {
    const size_t n = b.rec.size();
    threads = (int)std::max<size_t>(1, std::min<size_t>((size_t)threads, n / 4096 + 1));
    parts.assign((size_t)threads, std::string());
    std::vector<int64_t> cnt((size_t)threads, 0);
    auto work = [&](int t) {
        const size_t a = n * (size_t)t / (size_t)threads, z = n * (size_t)(t + 1) / (size_t)threads;
        std::string &out = parts[(size_t)t];
        out.reserve((z - a) * 112);
        for (size_t i = a; i < z; ++i) {
            if (b.first[i] < 0) continue;
            format_record(out, b.rec[i], res + b.first[i]);
            ++cnt[(size_t)t];
        }
    };
    std::vector<std::thread> th;
    for (int t = 1; t < threads; ++t) th.emplace_back(work, t);
    work(0);
    for (auto &x : th) x.join();
    for (int64_t k : cnt) *n_lines += k;
}

}  // namespace svtk

using namespace svtk;

// The VCF front end: the whole file is read into memory, cut into one piece per host thread at line boundaries,
// every piece parsed like thread_func parses a line (audit.c:65-173), the pieces' records and windows concatenated in file
// order.  tid_lo >= 0: only records on references tid_lo .. tid_hi (chrom - 1) are kept -- a region shard of a multi-GPU run.
static int vcf_front(const char *vcf_path, const svtk_params &params, int tid_lo, int tid_hi, unsigned flags, int n_ref, Batch &b, int *threads_used) {
    const bool quiet = (flags & SVTK_VCF_QUIET) != 0;
    const bool sharded = tid_lo >= 0;
    const bool owns_unplaced = !sharded || (flags & SVTK_VCF_OWNS_UNPLACED) != 0;
    // ---- read the file (pread by several threads: a 1 M-record VCF is ~150 MB)
    int fd = ::open(vcf_path, O_RDONLY);
    if (fd < 0) return fail(SVTK_ERR_IO, "cannot open %s: %s", vcf_path, strerror(errno));
    struct stat st;
    if (fstat(fd, &st) != 0) { ::close(fd); return fail(SVTK_ERR_IO, "cannot stat %s: %s", vcf_path, strerror(errno)); }
    const size_t size = (size_t)st.st_size;
    std::vector<char> text(size + 1);
    const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
    int T = (int)std::max<size_t>(1, std::min<size_t>(std::min(16u, hw), size / (1u << 20) + 1));
    if (const char *e = getenv("SVTK_VCF_THREADS")) if (*e) T = std::max(1, std::min(64, atoi(e)));     // tests: force the piece count
    *threads_used = T;
    {
        std::vector<int> ok((size_t)T, 1);
        auto rd = [&](int t) {
            size_t a = size * (size_t)t / (size_t)T;
            const size_t z = size * (size_t)(t + 1) / (size_t)T;
            while (a < z) {
                const ssize_t k = pread(fd, text.data() + a, z - a, (off_t)a);
                if (k <= 0) { ok[(size_t)t] = 0; return; }
                a += (size_t)k;
            }
        };
        std::vector<std::thread> th;
        for (int t = 1; t < T; ++t) th.emplace_back(rd, t);
        rd(0);
        for (auto &x : th) x.join();
        ::close(fd);
        for (int v : ok) if (!v) return fail(SVTK_ERR_IO, "cannot read %s", vcf_path);
    }
    text[size] = 0;

    // ---- parse: piece t = the lines that START in [cut[t], cut[t+1])
    std::vector<size_t> cut((size_t)T + 1, size);
    cut[0] = 0;
    for (int t = 1; t < T; ++t) {
        size_t p = size * (size_t)t / (size_t)T;
        p = std::max(p, cut[(size_t)t - 1]);
        const char *nl = p < size ? (const char *)memchr(text.data() + p, '\n', size - p) : nullptr;
        cut[(size_t)t] = nl ? (size_t)(nl - text.data()) + 1 : size;
    }
    struct Piece { Batch b; std::string err; };
    std::vector<Piece> piece((size_t)T);
    auto parse = [&](int t) {
        Piece &P = piece[(size_t)t];
        size_t p = cut[(size_t)t];
        const size_t end = cut[(size_t)t + 1];
        std::string msg;
        while (p < end) {
            const char *nl = (const char *)memchr(text.data() + p, '\n', size - p);
            const size_t e = nl ? (size_t)(nl - text.data()) + 1 : size;   // one past the line, newline included (fgets's view)
            size_t len = e - p;
            char *line = text.data() + p;
            p = e;
            if (len < 2 || line[0] == '#') continue;                        // audit.c:324
            if (line[len - 1] == '\n') line[--len] = 0;                     // audit.c:327-330 (a last line without one ends at text[size])
            Record r;
            r.chrom = 0;
            msg.clear();
            const int rc = parse_record(line, &r, &msg);
            // a record belongs to the shard that holds its reference; lines that never got that far, and records on references
            // no shard holds (they print NA), belong to the shard that owns the unplaced
            bool mine = !sharded;
            if (sharded) {
                if (rc == 0 && !msg.empty() && msg[0] == 'V') mine = owns_unplaced;            // "VCF: no index": no chromosome parsed
                else if (rc == 0 && msg.empty()) mine = false;                                 // silently skipped by every worker
                else if (r.chrom - 1 >= tid_lo && r.chrom - 1 <= tid_hi) mine = true;
                else if (r.chrom - 1 < 0 || (n_ref > 0 && r.chrom - 1 >= n_ref)) mine = owns_unplaced;
            }
            if (!mine) continue;
            if (!msg.empty()) P.err += msg;
            if (rc == 1) add_record(P.b, r, params);
        }
    };
    {
        std::vector<std::thread> th;
        for (int t = 1; t < T; ++t) th.emplace_back(parse, t);
        parse(0);
        for (auto &x : th) x.join();
    }
    // concatenate in file order (window indices shift by the windows of the pieces before); stderr messages in file order
    size_t nr = 0, nw = 0;
    for (auto &P : piece) { nr += P.b.rec.size(); nw += P.b.chrom.size(); }
    b.clear();
    b.rec.reserve(nr); b.first.reserve(nr); b.chrom.reserve(nw); b.start.reserve(nw); b.end.reserve(nw); b.pos.reserve(nw); b.kind.reserve(nw);
    for (auto &P : piece) {
        if (!P.err.empty() && !quiet) fwrite(P.err.data(), 1, P.err.size(), stderr);
        const int shift = (int)b.chrom.size();
        b.rec.insert(b.rec.end(), P.b.rec.begin(), P.b.rec.end());
        for (int f : P.b.first) b.first.push_back(f < 0 ? -1 : f + shift);
        b.chrom.insert(b.chrom.end(), P.b.chrom.begin(), P.b.chrom.end());
        b.start.insert(b.start.end(), P.b.start.begin(), P.b.start.end());
        b.end.insert(b.end.end(), P.b.end.begin(), P.b.end.end());
        b.pos.insert(b.pos.end(), P.b.pos.begin(), P.b.pos.end());
        b.kind.insert(b.kind.end(), P.b.kind.begin(), P.b.kind.end());
        P.b = Batch();
    }
    if (b.chrom.size() >= (size_t)0x7fffffff) return fail(SVTK_ERR_ARG, "more than 2^31 windows in one VCF: split the file");
    return 0;
}

// the stdout back end: the two [INFO] lines of process_vcf (audit.c:252,354) around one line per record that prints
static int vcf_back(const Batch &b, const int32_t *res, const char *out_path, unsigned flags, int threads, int64_t *n_lines, FILE **keep_open) {
    const bool no_info = (flags & SVTK_VCF_NO_INFO_LINES) != 0;
    const bool to_stdout = !out_path || !strcmp(out_path, "-");
    FILE *out = to_stdout ? stdout : fopen(out_path, "w");
    if (!out) return fail(SVTK_ERR_IO, "cannot open %s: %s", out_path, strerror(errno));
    if (!no_info) fputs("[INFO] Started processing variation file.\n", out);                 // audit.c:252
    const bool tee = (flags & SVTK_VCF_TEE_STDOUT) != 0 && !to_stdout;
    if (tee && !no_info) fputs("[INFO] Started processing variation file.\n", stdout);
    std::vector<std::string> parts;
    format_all(b, res, threads, parts, n_lines);
    for (auto &s : parts) { fwrite(s.data(), 1, s.size(), out); if (tee) fwrite(s.data(), 1, s.size(), stdout); }
    *keep_open = out;
    return 0;
}

static int vcf_back_close(FILE *out, const char *out_path, unsigned flags, bool ok) {
    const bool to_stdout = !out_path || !strcmp(out_path, "-");
    if (ok && !(flags & SVTK_VCF_NO_INFO_LINES)) {
        fputs("[INFO] Ended processing variation file\n", out);   // audit.c:354
        if ((flags & SVTK_VCF_TEE_STDOUT) && !to_stdout) fputs("[INFO] Ended processing variation file\n", stdout);
    }
    if ((flags & SVTK_VCF_TEE_STDOUT) && !to_stdout) fflush(stdout);
    if (to_stdout) fflush(out); else fclose(out);
    return 0;
}

static int process_vcf_impl(svtk_ctx *ctx, const char *vcf_path, const char *out_path, const char *fasta_path, int tol, int64_t *n_records,
                            int64_t *n_consensus, int tid_lo = -1, int tid_hi = -1, unsigned flags = 0) {
    Ctx *c = reinterpret_cast<Ctx *>(ctx);
    if (!c || !vcf_path) return fail(SVTK_ERR_ARG, "null argument");
    if (fasta_path && !c->reads_from_bam) return fail(SVTK_ERR_STATE, "insertion consensus needs the BAM image: load the reads with svtk_load_bam");
    if (!c->have_reads) return fail(SVTK_ERR_STATE, "no reads loaded (svtk_load_reads)");
    Batch b;
    int T = 1;
    if (int rc = vcf_front(vcf_path, c->params, tid_lo, tid_hi, flags, c->reads.n_tid, b, &T)) return rc;

    // ---- refine (ONE batch), write the record lines, then the consensus side file
    FILE *fa = fasta_path ? fopen(fasta_path, "w") : nullptr;
    if (fasta_path && !fa) return fail(SVTK_ERR_IO, "cannot open %s: %s", fasta_path, strerror(errno));
    struct Closer { FILE *f; ~Closer() { if (f) fclose(f); } } fa_closer{fa};
    int64_t n_lines = 0, n_cons = 0;
    std::vector<int32_t> res(std::max<size_t>(b.chrom.size(), 1));
    const svtk_windows w = {(int32_t)b.chrom.size(), b.chrom.data(), b.start.data(), b.end.data(), b.pos.data(), b.kind.data()};
    int rc = 0;
    if (w.n_win > 0) rc = svtk_refine_windows(ctx, &w, res.data());
    if (rc) return rc;
    FILE *out = nullptr;
    if (int rc2 = vcf_back(b, res.data(), out_path, flags, T, &n_lines, &out)) return rc2;
    // the record lines do not depend on the consensus side file: they are out already, and a read set that fails its POA is
    // skipped (and reported) instead of failing the batch
    if (fa && w.n_win > 0) rc = write_ins_consensus(ctx, b, w, res.data(), tol, fa, &n_cons);
    vcf_back_close(out, out_path, flags, rc == 0);
    if (n_records) *n_records = n_lines;
    if (n_consensus) *n_consensus = n_cons;
    return rc;
}

// ---- the front and back ends alone (no GPU, no ctx): what windows process_vcf builds for a VCF, and the lines it prints for given
// window results.  Parity aids for the host logic (tests run them against the reference's captured stdout on CPU-only boxes).
extern "C" int svtk_vcf_windows(const char *vcf_path, const svtk_params *params, int32_t n_ref, int32_t tid_lo, int32_t tid_hi, uint32_t flags,
                                int64_t cap, int32_t *chrom, uint32_t *start, uint32_t *end, uint32_t *pos, uint8_t *kind, int64_t *n_win) {
    if (!vcf_path || !n_win) return fail(SVTK_ERR_ARG, "null argument");
    const svtk_params def = {20000, 10000, 2000, 500, 5, 3};
    Batch b;
    int T = 1;
    if (int rc = vcf_front(vcf_path, params ? *params : def, tid_lo, tid_hi, flags, n_ref, b, &T)) return rc;
    *n_win = (int64_t)b.chrom.size();
    const size_t k = (size_t)std::min<int64_t>(cap, *n_win);
    if (k && (!chrom || !start || !end || !pos || !kind)) return fail(SVTK_ERR_ARG, "null window array");
    if (k) {
        memcpy(chrom, b.chrom.data(), 4 * k); memcpy(start, b.start.data(), 4 * k); memcpy(end, b.end.data(), 4 * k);
        memcpy(pos, b.pos.data(), 4 * k); memcpy(kind, b.kind.data(), k);
    }
    return 0;
}

extern "C" int svtk_vcf_format(const char *vcf_path, const svtk_params *params, int32_t n_ref, int32_t tid_lo, int32_t tid_hi, uint32_t flags,
                               const int32_t *results, int64_t n_results, const char *out_path, int64_t *n_records) {
    if (!vcf_path) return fail(SVTK_ERR_ARG, "null argument");
    const svtk_params def = {20000, 10000, 2000, 500, 5, 3};
    Batch b;
    int T = 1;
    if (int rc = vcf_front(vcf_path, params ? *params : def, tid_lo, tid_hi, flags, n_ref, b, &T)) return rc;
    if ((int64_t)b.chrom.size() != n_results) return fail(SVTK_ERR_ARG, "the VCF builds %zu windows, %lld results given", b.chrom.size(), (long long)n_results);
    if (n_results > 0 && !results) return fail(SVTK_ERR_ARG, "null results");
    int64_t n_lines = 0;
    FILE *out = nullptr;
    const int32_t dummy = -1;
    if (int rc = vcf_back(b, n_results ? results : &dummy, out_path, flags, T, &n_lines, &out)) return rc;
    vcf_back_close(out, out_path, flags, true);
    if (n_records) *n_records = n_lines;
    return 0;
}

extern "C" int svtk_process_vcf(svtk_ctx *ctx, const char *vcf_path, const char *out_path, int64_t *n_records) {
    return process_vcf_impl(ctx, vcf_path, out_path, nullptr, 0, n_records, nullptr);
}

extern "C" int svtk_process_vcf_flags(svtk_ctx *ctx, const char *vcf_path, const char *out_path, const char *fasta_path, int32_t tol, uint32_t flags,
                                      int64_t *n_records, int64_t *n_consensus) {
    if (tol < 0) tol = reinterpret_cast<Ctx *>(ctx) ? reinterpret_cast<Ctx *>(ctx)->params.consensus_interval : 0;
    return process_vcf_impl(ctx, vcf_path, out_path, fasta_path, tol, n_records, n_consensus, -1, -1, flags);
}

// One region shard of a whole job, file to file: the VCF's records are cut into n_shards contiguous runs (VCF order == genomic
// order) of equal estimated work -- compressed BAM bytes under each record's windows, by the .bai's linear index -- and shard
// `shard` loads only what ITS windows can see (svtk_load_bam_windows: cuts fall inside references), refines them and writes its
// lines.  Every shard computes the same cut points; shard outputs concatenated in order are the whole output.
extern "C" int svtk_audit_shard(svtk_ctx *ctx, const char *bam_path, const char *bai_path, const char *vcf_path, const char *out_path, int32_t shard,
                                int32_t n_shards, uint32_t flags, int64_t *n_records, svtk_bam_stats *stats) {
    Ctx *c = reinterpret_cast<Ctx *>(ctx);
    if (!c || !bam_path || !vcf_path) return fail(SVTK_ERR_ARG, "null argument");
    if (n_shards < 1 || shard < 0 || shard >= n_shards) return fail(SVTK_ERR_ARG, "bad shard %d of %d", shard, n_shards);
    Batch b;
    int T = 1;
    const unsigned fl = flags | (shard == 0 ? 0u : (unsigned)SVTK_VCF_QUIET);          // malformed lines are reported once, by shard 0
    if (int rc = vcf_front(vcf_path, c->params, -1, -1, fl, 0, b, &T)) return rc;
    const size_t nr = b.rec.size(), nw = b.chrom.size();
    // work per record = compressed bytes under its windows
    std::vector<double> ww(std::max<size_t>(nw, 1), 1.0);
    if (nw) if (int rc = bam_window_weights(bam_path, bai_path, (int)nw, b.chrom.data(), b.start.data(), b.end.data(), ww.data())) return rc;
    std::vector<double> cum(nr + 1, 0.0);
    std::vector<int> w_first(nr + 1, (int)nw);                 // first window at or after record i
    {
        int next = (int)nw;
        for (size_t i = nr; i-- > 0;) { if (b.first[i] >= 0) next = b.first[i]; w_first[i] = next; }
    }
    for (size_t i = 0; i < nr; ++i) {
        double w = 1.0;
        for (int k = w_first[i]; k < w_first[i + 1]; ++k) w += ww[(size_t)k];
        cum[i + 1] = cum[i] + w;
    }
    auto cut = [&](int s) -> size_t {
        if (s <= 0) return 0;
        if (s >= n_shards) return nr;
        const double target = cum[nr] * (double)s / (double)n_shards;
        return (size_t)(std::lower_bound(cum.begin(), cum.end(), target) - cum.begin());
    };
    const size_t r0 = std::min(cut(shard), nr), r1 = std::min(std::max(cut(shard + 1), r0), nr);
    const int w0 = w_first[r0], w1 = w_first[r1];
    // the shard's own batch (window indices rebased)
    Batch m;
    for (size_t i = r0; i < r1; ++i) { m.rec.push_back(b.rec[i]); m.first.push_back(b.first[i] < 0 ? -1 : b.first[i] - w0); }
    m.chrom.assign(b.chrom.begin() + w0, b.chrom.begin() + w1); m.start.assign(b.start.begin() + w0, b.start.begin() + w1);
    m.end.assign(b.end.begin() + w0, b.end.begin() + w1); m.pos.assign(b.pos.begin() + w0, b.pos.begin() + w1);
    m.kind.assign(b.kind.begin() + w0, b.kind.begin() + w1);
    const svtk_windows w = {(int32_t)m.chrom.size(), m.chrom.data(), m.start.data(), m.end.data(), m.pos.data(), m.kind.data()};
    if (int rc = svtk_load_bam_windows(ctx, bam_path, bai_path, &w, stats)) return rc;
    std::vector<int32_t> res(std::max<size_t>(m.chrom.size(), 1));
    if (w.n_win > 0) if (int rc = svtk_refine_windows(ctx, &w, res.data())) return rc;
    int64_t n_lines = 0;
    FILE *out = nullptr;
    if (int rc = vcf_back(m, res.data(), out_path, flags, T, &n_lines, &out)) return rc;
    vcf_back_close(out, out_path, flags, true);
    if (n_records) *n_records = n_lines;
    return 0;
}

extern "C" int svtk_process_vcf_region(svtk_ctx *ctx, const char *vcf_path, const char *out_path, int32_t tid_lo, int32_t tid_hi, uint32_t flags,
                                       int64_t *n_records) {
    if (tid_lo < 0 || tid_hi < tid_lo) return fail(SVTK_ERR_ARG, "bad reference range");
    return process_vcf_impl(ctx, vcf_path, out_path, nullptr, 0, n_records, nullptr, tid_lo, tid_hi, flags);
}

extern "C" int svtk_process_vcf_ins(svtk_ctx *ctx, const char *vcf_path, const char *out_path, const char *fasta_path, int32_t tol, int64_t *n_records,
                                    int64_t *n_consensus) {
    if (!fasta_path) return fail(SVTK_ERR_ARG, "null fasta path");
    if (tol < 0) tol = reinterpret_cast<Ctx *>(ctx) ? reinterpret_cast<Ctx *>(ctx)->params.consensus_interval : 0;
    return process_vcf_impl(ctx, vcf_path, out_path, fasta_path, tol, n_records, n_consensus);
}
