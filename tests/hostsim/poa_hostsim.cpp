// TEST BUILD ONLY: svtk_poa_core.cuh compiled as plain C++ with warp width 1 (POA_HOSTSIM), so the K3 algorithm
// can be checked against the abPOA golden vectors on a machine without a GPU.  Never linked into libsvtk_b200.so.
#define POA_HOSTSIM 1
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <vector>
struct int2 { int x, y; };
struct int4 { int x, y, z, w; };
static inline int4 make_int4(int x, int y, int z, int w) { int4 r; r.x = x; r.y = y; r.z = z; r.w = w; return r; }
#include "../../svtrek_b200/csrc/svtk_poa_core.cuh"

using namespace svtk_poa;

extern "C" int hostsim_poa(int n, const long long *off, const uint8_t *bases, uint8_t *cons, int cons_cap, int32_t *cov,
                           int32_t *best_score, int32_t *node_n, int32_t *n_ops, uint64_t *ops_hash, long long *cells,
                           int dump_read, int32_t *dump, int dump_cap, int *status, long long *cells_total) {
    long long total = off[n] - off[0];
    int maxlen = 0;
    for (int i = 0; i < n; ++i) if (off[i + 1] - off[i] > maxlen) maxlen = (int)(off[i + 1] - off[i]);
    PoaState s;
    memset(&s, 0, sizeof(s));
    s.node_cap = (int)total + 2;
    s.edge_cap = 16 * s.node_cap + 64;
    s.cigar_cap = s.node_cap + maxlen + 8;
    long long want = (long long)s.node_cap * (long long)(maxlen + 1) * 20;
    if (want > (3ll << 30)) want = 3ll << 30;
    s.cell_bytes = want;
    std::vector<uint8_t> base(s.node_cap);
    std::vector<int> ints((size_t)s.node_cap * 24 + 16), eid(s.edge_cap), ew(s.edge_cap);
    std::vector<long long> roff(s.node_cap);
    std::vector<unsigned char> cellbuf((size_t)s.cell_bytes);
    std::vector<int2> cig(s.cigar_cap);
    std::vector<int4> rmeta(s.node_cap), rout(s.node_cap), rinfo(s.node_cap);
    std::vector<int> eidx(s.edge_cap);
    int *p = ints.data();
    auto take = [&](int k) { int *r = p; p += (size_t)k * s.node_cap; return r; };
    s.base = base.data();
    s.in_off = take(1); s.in_n = take(1); s.in_cap = take(1); s.out_off = take(1); s.out_n = take(1); s.out_cap = take(1);
    s.aln_n = take(1); s.aln_id = take(ALN_MAX); s.n_read = take(1);
    s.idx2id = take(1); s.id2idx = take(1); s.max_left = take(1); s.max_right = take(1); s.remain = take(1);
    s.deg = take(1); s.queue = take(1); s.dp_beg = take(1); s.dp_end = take(1);
    s.edge_id = eid.data(); s.edge_w = ew.data(); s.row_off = roff.data(); s.cells = cellbuf.data(); s.cigar = cig.data();
    s.rowmeta = rmeta.data(); s.rowout = rout.data(); s.rowinfo = rinfo.data(); s.edge_idx = eidx.data();
    PoaTrace tr;
    tr.best_score = best_score; tr.node_n = node_n; tr.n_ops = n_ops; tr.ops_hash = ops_hash; tr.cells = cells;
    tr.dump_read = dump_read; tr.dump = dump; tr.dump_cap = dump_cap;
    PoaXchg xch;
    int len = poa_run_set<1>(s, xch, 0, n, off, bases, cons, cov, cons_cap, best_score ? &tr : nullptr);
    *status = s.status;
    if (cells_total) *cells_total = s.cells_done;
    return len;
}
