// svtk_poa_core.cuh -- K3: insertion consensus by partial-order alignment, equivalent to abPOA's default
// abpoa_msa() path (reference submodule abPOA/ v1.5.4):
//     abpoa_msa                abpoa_align.c:401        -> abpoa_poa :312  (no seeding, reads in input order)
//     alignment                abpoa_align_simd.c:1250  convex gap (4,2 / 24,1), match 2, mismatch 4, global,
//                              adaptive band w = 10 + (int)(0.01f * qlen)        (:474, abpoa_align.h:34-35)
//         first row            :669-688     row DP :935-1074     row arg-max -> band hints :1107-1130
//         best end cell        :1092-1105   backtrack :309-458 (M -> E1 -> E2 -> F1 -> F2 -> M, cur_op machine)
//     graph fusion             abpoa_graph.c:688-773    add node/edge :477-553, aligned groups :450-465
//     topological sort         abpoa_graph.c:322        BFS index :221-266, exchange sort of edges :192-219,
//                              max_remain reverse BFS :268-309
//     consensus                abpoa_output.c:477-547   heaviest bundling, walk from SRC :375-391
//
// Execution model: one CTA owns one read set -- a single warp (NW = 1, the throughput shape: up to 16 sets per SM
// in flight) or, for the few heaviest sets of a batch, 8 warps (NW = 8, poa_align_dp_cta).  The row DP is column
// parallel across the threads -- one cell of the five-state convex recurrence per thread and step, the two
// in-row gap states (F1/F2) by a shuffle prefix-max (plus a shared-memory exchange of warp totals at NW = 8), row
// arg-max by warp reductions; edge sorting, max_remain (pointer jumping) and the DP's row descriptions are thread
// parallel; the pointer-chasing phases (backtrack, graph fusion, the BFS index, bundling) run on thread 0 while
// the other warps of the SM are in their DP phase.  abPOA's SIMD vectors are a storage detail of the CPU code: this file implements
// the vector-width-1 semantics (cells outside a row's band read as inf_min), which is what every real-valued
// cell of abPOA computes; SURVEY 7.3-6 measured the four CPU vector widths to agree on the same inputs.
//
// The same source compiles as plain C++ with POA_HOSTSIM defined (warp width 1): tests/ build that variant
// to exercise the algorithm against the abPOA golden vectors on machines without a GPU.  It is a test build
// of this file, not a product path: libsvtk_b200.so contains only the device instantiation.
#pragma once
#include <stdint.h>

#ifdef POA_HOSTSIM
#define PD static inline
#define POA_NL 1
static inline int poa_lane() { return 0; }
static inline void poa_sync() {}
static inline int poa_shfl(int v, int) { return v; }
static inline long long poa_shfl_ll(long long v, int) { return v; }
static inline int poa_shfl_up(int v, int) { return v; }
static inline int poa_max_all(int v) { return v; }
static inline int poa_min_all(int v) { return v; }
static inline bool poa_any(bool v) { return v; }
static inline void poa_touch(const void *) {}
#else
#define PD __device__ __forceinline__
#define POA_NL 32
PD int poa_lane() { return threadIdx.x & 31; }
PD void poa_sync() { __syncwarp(); }
PD int poa_shfl(int v, int src) { return __shfl_sync(0xffffffffu, v, src); }
PD long long poa_shfl_ll(long long v, int src) { return __shfl_sync(0xffffffffu, v, src); }
PD int poa_shfl_up(int v, int d) { return __shfl_up_sync(0xffffffffu, v, d); }
PD int poa_max_all(int v) { return __reduce_max_sync(0xffffffffu, v); }
PD int poa_min_all(int v) { return __reduce_min_sync(0xffffffffu, v); }
PD bool poa_any(bool v) { return __any_sync(0xffffffffu, v) != 0; }
// request a line into L1 without consuming it (the value is never read, so nothing waits for it)
PD void poa_touch(const void *p) { unsigned v; asm volatile("ld.global.ca.u32 %0, [%1];" : "=r"(v) : "l"(p)); (void)v; }
#endif
#if defined(POA_HOSTSIM)
static inline long long poa_clock() { return 0; }
template <int NW> static inline int poa_tid() { return 0; }
template <int NW> static inline void poa_bar() {}
#else
PD long long poa_clock() { return clock64(); }
// a read set is owned by NW warps (NW = 1: one warp, warp-level sync; NW > 1: the whole CTA, block barriers)
template <int NW> PD int poa_tid() { return NW == 1 ? (int)(threadIdx.x & 31) : (int)threadIdx.x; }
template <int NW> PD void poa_bar() { if (NW == 1) __syncwarp(); else __syncthreads(); }
#endif

namespace svtk_poa {

// abPOA defaults: abpoa_align.h:9-14, abpoa.h:16-17, abpoa_align.c:100-157
constexpr int MATCH = 2, MISMATCH = 4, GAP_O1 = 4, GAP_E1 = 2, GAP_O2 = 24, GAP_E2 = 1;
constexpr int GAP_OE1 = GAP_O1 + GAP_E1, GAP_OE2 = GAP_O2 + GAP_E2;
constexpr int EXTRA_B = 10;
constexpr float EXTRA_F = 0.01f;
constexpr int SRC = 0, SINK = 1;                              // abpoa.h:31-32
constexpr int OP_M = 1, OP_E1 = 2, OP_E2 = 4, OP_E = 6, OP_F1 = 8, OP_F2 = 16, OP_F = 24, OP_ALL = 31;   // abpoa_align.h:20-27
constexpr int CG_MATCH = 0, CG_INS = 1, CG_DEL = 2;           // abpoa.h:20-22
#ifndef POA_ROW_CB
#define POA_ROW_CB 2
#endif
constexpr int ROW_CB = POA_ROW_CB;                                     // 32-column chunks of a row whose loads are issued together
#ifndef POA_PRED_REG
#define POA_PRED_REG 3
#endif
constexpr int PRED_REG = POA_PRED_REG;                                   // predecessors of a row held in registers (most nodes have 1-2 in-edges)
constexpr int NPL = 3;                                        // score planes kept per DP cell: H, E1, E2 (F1 / F2 are recomputed by the backtrack)
constexpr int ALN_MAX = 4;                                    // an aligned group holds distinct bases: <= 5 nodes (m = 5)

enum { POA_OK = 0, POA_ERR_NODES = 1, POA_ERR_EDGES = 2, POA_ERR_CELLS = 3, POA_ERR_CIGAR = 4, POA_ERR_BACKTRACK = 5,
       POA_ERR_BAND = 6, POA_ERR_BFS = 7, POA_ERR_CONS = 8, POA_ERR_INPUT = 9 };

// inf_min of the two score widths (abpoa_align_simd.c:1293-1299)
constexpr int INF_MIN16 = -32768 + 25 + 512 * 2;
constexpr int INF_MIN32 = (-2147483647 - 1) + 25 + 512 * 2;

struct PoaState {
    // capacities of this worker's scratch
    int node_cap, edge_cap, cigar_cap;
    long long cell_bytes;
    // graph (abpoa_node_t / abpoa_graph_t, abpoa.h:89-110) as SoA
    int node_n, edge_top;
    uint8_t *base;
    int *in_off, *in_n, *in_cap, *out_off, *out_n, *out_cap;   // edge lists live in the pool below
    int *edge_id, *edge_w;
    int *aln_n, *aln_id;                                       // aligned_node_id, ALN_MAX per node
    int *n_read;
    int *idx2id, *id2idx, *max_left, *max_right, *remain;      // index_to_node_id, node_id_to_index, max_pos_left/right, max_remain
    int *deg, *queue;                                          // BFS scratch
    // what the row DP streams, rebuilt by every topological sort: per INDEX {node id, base | in-degree << 8, first
    // in-edge slot, max_remain}, and per in-edge slot the predecessor's index
    int4 *rowmeta;
    int4 *rowinfo;                                             // per INDEX {dp_beg, dp_end, row_off low, row_off high}: one load for the backtrack
    int4 *rowout;                                              // per INDEX {out-degree, first out-edge slot, target of out-edge 0, of out-edge 1}
    int *edge_idx;
    // one alignment
    int *dp_beg, *dp_end;
    long long *row_off;                                        // first cell of each row's band (in cells)
    unsigned char *cells;                                      // rows x band x {H, E1, E2, F1, F2}, int16 or int32
    int2 *cigar;
    int n_cigar;
    int best_score, best_i, best_j, ret_len;
    long long cells_peak;                                      // most band cells any one alignment of this set used (sizing statistics)
    long long cells_done, cells_full;                          // band cells / full-matrix-equivalent cells of this read set
    long long t_phase[4];                                      // SM cycles of this read set in: row DP, backtrack, graph fusion, topological sort
    int status;
};

PD int poa_score(int node_base, int q) { return (node_base > 3 || q > 3) ? 0 : (node_base == q ? MATCH : -MISMATCH); }   // gen_simple_mat, abpoa_align.c:12-25
PD int imax(int a, int b) { return a > b ? a : b; }
PD int imin(int a, int b) { return a < b ? a : b; }

// ----------------------------------------------------------------------------------------------- graph (lane 0)
PD int poa_add_node(PoaState &s, uint8_t b) {                 // abpoa_add_graph_node, abpoa_graph.c:468
    if (s.node_n >= s.node_cap) { s.status = POA_ERR_NODES; return -1; }
    const int id = s.node_n++;
    s.base[id] = b;
    s.in_n[id] = s.out_n[id] = s.in_cap[id] = s.out_cap[id] = s.in_off[id] = s.out_off[id] = 0;
    s.aln_n[id] = 0; s.n_read[id] = 0;
    return id;
}

PD bool poa_grow(PoaState &s, int *off, int *n, int *cap, int id) {
    if (n[id] < cap[id]) return true;
    const int nc = cap[id] ? cap[id] * 2 : 2;
    if (s.edge_top + nc > s.edge_cap) { s.status = POA_ERR_EDGES; return false; }
    const int no = s.edge_top;
    s.edge_top += nc;
    for (int k = 0; k < n[id]; ++k) { s.edge_id[no + k] = s.edge_id[off[id] + k]; s.edge_w[no + k] = s.edge_w[off[id] + k]; }
    off[id] = no; cap[id] = nc;
    return true;
}

PD void poa_add_edge(PoaState &s, int from, int to, int check, int w) {   // abpoa_add_graph_edge, abpoa_graph.c:477
    bool exist = false;
    if (check) {
        for (int i = 0; i < s.in_n[to]; ++i)
            if (s.edge_id[s.in_off[to] + i] == from) { s.edge_w[s.in_off[to] + i] += w; break; }
        for (int i = 0; i < s.out_n[from]; ++i)
            if (s.edge_id[s.out_off[from] + i] == to) { s.edge_w[s.out_off[from] + i] += w; exist = true; break; }
    }
    if (!exist) {
        if (!poa_grow(s, s.in_off, s.in_n, s.in_cap, to)) return;
        s.edge_id[s.in_off[to] + s.in_n[to]] = from; s.edge_w[s.in_off[to] + s.in_n[to]] = w; ++s.in_n[to];
        if (!poa_grow(s, s.out_off, s.out_n, s.out_cap, from)) return;
        s.edge_id[s.out_off[from] + s.out_n[from]] = to; s.edge_w[s.out_off[from] + s.out_n[from]] = w; ++s.out_n[from];
    }
    s.n_read[from] += 1;
}

PD void poa_aln_push(PoaState &s, int node, int other) {
    if (s.aln_n[node] < ALN_MAX) s.aln_id[node * ALN_MAX + s.aln_n[node]++] = other;
    else s.status = POA_ERR_NODES;
}

PD void poa_add_aligned(PoaState &s, int node_id, int new_id) {            // abpoa_add_graph_aligned_node, abpoa_graph.c:450
    const int n = s.aln_n[node_id];
    for (int i = 0; i < n; ++i) {
        const int a = s.aln_id[node_id * ALN_MAX + i];
        poa_aln_push(s, a, new_id);
        poa_aln_push(s, new_id, a);
    }
    poa_aln_push(s, node_id, new_id);
    poa_aln_push(s, new_id, node_id);
}

// BFS index (abpoa_BFS_set_node_index, abpoa_graph.c:221): FIFO Kahn order, a node enters the queue together with its
// aligned group once every member has in-degree 0.  Uses the out-edge order BEFORE this round's edge sort.
// One thread, one chain of dependent loads per node; the caller has laid out what a step needs so that the chain is three loads
// deep instead of five: rowout[id] (free until the end of the sort) = {out-degree, first out-edge slot, targets of edges 0 and 1},
// deg[id] = remaining in-degree | aligned-group size << 24.
PD void poa_bfs_index(PoaState &s) {
    int head = 0, tail = 0, index = 0;
    s.queue[tail++] = SRC;
    while (head < tail) {
        const int cur = s.queue[head++];
        s.idx2id[index] = cur; s.id2idx[cur] = index++;
        if (cur == SINK) { if (index != s.node_n) s.status = POA_ERR_BFS; return; }
        const int4 out = s.rowout[cur];
        const int no = out.x, oo = out.y;
        for (int i = 0; i < no; ++i) {
            const int o = i == 0 ? out.z : (i == 1 ? out.w : s.edge_id[oo + i]);
            const int d = s.deg[o] - 1;
            s.deg[o] = d;
            if ((d & 0xffffff) == 0) {
                bool ok = true;
                const int na = (int)((unsigned)d >> 24);
                for (int j = 0; j < na; ++j) if ((s.deg[s.aln_id[o * ALN_MAX + j]] & 0xffffff) != 0) { ok = false; break; }
                if (!ok) continue;
                if (tail + 1 + na > s.node_cap) { s.status = POA_ERR_BFS; return; }
                s.queue[tail++] = o;
                for (int j = 0; j < na; ++j) s.queue[tail++] = s.aln_id[o * ALN_MAX + j];
            }
        }
    }
    s.status = POA_ERR_BFS;
}

// descending exchange sort of one edge list, swap sequence as in abpoa_sort_in_out_ids (abpoa_graph.c:192-219):
// not stable, and applied to the order the previous rounds left behind -- tie order decides the backtrack's choices
PD void poa_sort_edges(int *id, int *w, int n) {
    for (int j = 0; j < n - 1; ++j)
        for (int k = j + 1; k < n; ++k)
            if (w[j] < w[k]) { int t = id[j]; id[j] = id[k]; id[k] = t; t = w[j]; w[j] = w[k]; w[k] = t; }
}

// abpoa_topological_sort (abpoa_graph.c:322): all lanes enter
template <int NW>
PD void poa_topological_sort(PoaState &sh) {
    const int lane = poa_tid<NW>();
    poa_bar<NW>();
    PoaState s = sh;                                           // register copy: the shared struct would be re-read after every store
    const int n = s.node_n;
    for (int i = lane; i < n; i += (NW * POA_NL)) {
        const int on = s.out_n[i], oo = s.out_off[i];
        s.deg[i] = s.in_n[i] | (s.aln_n[i] << 24);             // in-degree < 2^24 (node_cap), group size <= ALN_MAX
        s.rowout[i] = make_int4(on, oo, on > 0 ? s.edge_id[oo] : -1, on > 1 ? s.edge_id[oo + 1] : -1);
    }
    poa_bar<NW>();
    if (lane == 0) { poa_bfs_index(s); sh.status = s.status; }
    poa_bar<NW>();
    if (sh.status) return;
    constexpr int NT = NW * POA_NL;
    // max_remain (abpoa_BFS_set_node_remain, abpoa_graph.c:268): remain[v] = remain[heaviest successor of v] + 1 with
    // remain[SINK] = -1, the heaviest successor being the FIRST out-edge once the list is sorted by weight (the
    // reference takes the first strict maximum).  The reference walks a reverse BFS; the values only depend on the
    // successor chain, so they are computed here by pointer jumping (ceil(log2 n) rounds, all threads) instead of
    // one thread chasing n pointers.  Buffers: (deg, remain) <-> (queue, max_left).
    for (int i = lane; i < n; i += NT) {                       // independent per node: threads take nodes
        poa_sort_edges(s.edge_id + s.in_off[i], s.edge_w + s.in_off[i], s.in_n[i]);
        poa_sort_edges(s.edge_id + s.out_off[i], s.edge_w + s.out_off[i], s.out_n[i]);
        int nx = SINK, d = 0;
        if (i != SINK) {
            if (s.out_n[i] <= 0) sh.status = POA_ERR_BFS;      // every node reaches SINK in a well-formed graph
            else { nx = s.edge_id[s.out_off[i]]; d = 1; }
        }
        s.deg[i] = nx; s.remain[i] = d;
    }
    poa_bar<NW>();
    if (sh.status) return;
    {
        int *nxa = s.deg, *da = s.remain, *nxb = s.queue, *db = s.max_left;
        for (int span = 1; span < n; span <<= 1) {
            for (int i = lane; i < n; i += NT) {
                const int nx = nxa[i];
                db[i] = da[i] + da[nx];
                nxb[i] = nxa[nx];
            }
            poa_bar<NW>();
            int *t = nxa; nxa = nxb; nxb = t;
            t = da; da = db; db = t;
        }
        if (da != s.remain) {
            for (int i = lane; i < n; i += NT) s.remain[i] = da[i];
            poa_bar<NW>();
        }
    }
    for (int i = lane; i < n; i += NT) {
        s.remain[i] -= 1;
        s.max_right[i] = 0; s.max_left[i] = n;                 // abpoa_graph.c:344-347
    }
    poa_bar<NW>();
    for (int i = lane; i < n; i += (NW * POA_NL)) {                   // the DP's streaming view of the sorted graph
        const int id = s.idx2id[i];
        const int pn = s.in_n[id], po = s.in_off[id];
        int4 m;
        m.x = id; m.y = (int)s.base[id] | (pn << 8); m.z = po; m.w = s.remain[id];
        s.rowmeta[i] = m;
        for (int k = 0; k < pn; ++k) s.edge_idx[po + k] = s.id2idx[s.edge_id[po + k]];
        const int on = s.out_n[id], oo = s.out_off[id];
        s.rowout[i] = make_int4(on, oo, on > 0 ? s.edge_id[oo] : -1, on > 1 ? s.edge_id[oo + 1] : -1);
    }
    poa_bar<NW>();
}

// first read: abpoa_add_graph_sequence (abpoa_graph.c:555).  lane 0
PD void poa_add_first(PoaState &s, const uint8_t *q, int qlen) {
    int last = SRC;
    for (int i = 0; i < qlen; ++i) {
        const int cur = poa_add_node(s, q[i]);
        if (cur < 0) return;
        poa_add_edge(s, last, cur, 0, 1);
        if (s.status) return;
        last = cur;
    }
    poa_add_edge(s, last, SINK, 0, 1);
}

// abpoa_add_subgraph_alignment (abpoa_graph.c:688) with beg = SRC, end = SINK, inc_both_ends = 1, all weights 1.
// The cigar is in backtrack order (last op first); ops are consumed from the back.
// All lanes of the first warp enter, lane 0 does the work (one thread chasing pointers: every op reads what the previous one
// wrote).  Measured, not kept: the other lanes requesting the node and edge-list lines of the next 32 ops into L1 ahead of lane 0
// (the cigar says which) -- fusion time unchanged (+5 %): the chain is not waiting on those lines.
PD void poa_add_alignment(PoaState &s, const uint8_t *q, int qlen, int lane) {
    if (s.n_cigar == 0) return;                                // abpoa_graph.c:706-709
    int query_id = -1, last_new = 0, last_id = SRC;
    for (int c0 = s.n_cigar - 1; c0 >= 0; c0 -= POA_NL) {
        bool stop = false;
        if (lane == 0)
        for (int c = c0; c > c0 - POA_NL && c >= 0; --c) {
            const int op = s.cigar[c].x, arg = s.cigar[c].y;
            if (op == CG_MATCH) {
                const int node_id = arg;
                ++query_id;
                const uint8_t b = q[query_id];
                if (s.base[node_id] != b) {                    // mismatch: reuse an aligned node with this base, else a new one
                    int aligned = -1;
                    for (int i = 0; i < s.aln_n[node_id]; ++i) {
                        const int a = s.aln_id[node_id * ALN_MAX + i];
                        if (s.base[a] == b) { aligned = a; break; }
                    }
                    if (aligned != -1) {
                        poa_add_edge(s, last_id, aligned, 1 - last_new, 1);
                        last_id = aligned; last_new = 0;
                    } else {
                        const int nid = poa_add_node(s, b);
                        if (nid < 0) { stop = true; break; }
                        poa_add_edge(s, last_id, nid, 0, 1);
                        last_id = nid; last_new = 1;
                        poa_add_aligned(s, node_id, nid);
                    }
                } else {
                    poa_add_edge(s, last_id, node_id, 1 - last_new, 1);
                    last_id = node_id; last_new = 0;
                }
            } else if (op == CG_INS) {
                for (int l = 0; l < arg; ++l) {
                    ++query_id;
                    const int nid = poa_add_node(s, q[query_id]);
                    if (nid < 0) { stop = true; break; }
                    poa_add_edge(s, last_id, nid, 0, 1);
                    last_id = nid; last_new = 1;
                    if (s.status) { stop = true; break; }
                }
            }                                                  // CG_DEL: nothing
            if (stop || s.status) { stop = true; break; }
        }
        if (poa_any(stop)) return;                             // out of nodes / edges: the caller sees s.status (lane 0's copy)
    }
    (void)qlen;
    if (lane == 0) poa_add_edge(s, last_id, SINK, 1 - last_new, 1);
}

// ----------------------------------------------------------------------------------------------- alignment
template <typename SC>
PD int poa_cell(const PoaState &s, int i, int plane, int j, int inf_min) {
    const int b = s.dp_beg[i], e = s.dp_end[i];
    if (j < b || j > e) return inf_min;
    const int w = e - b + 1;
    return (int)reinterpret_cast<const SC *>(s.cells)[s.row_off[i] * NPL + (long long)plane * w + (j - b)];
}

PD void poa_push_cigar(PoaState &s, int op, int arg) {
    if (s.n_cigar >= s.cigar_cap) { s.status = POA_ERR_CIGAR; return; }
    s.cigar[s.n_cigar].x = op; s.cigar[s.n_cigar].y = arg; ++s.n_cigar;
}

// One DP row as the backtrack sees it: band and cells from ONE 16-byte load (rowinfo).
template <typename SC>
struct PoaRow {
    int b, e, w;
    const SC *c;
};
template <typename SC>
PD PoaRow<SC> poa_row(const PoaState &s, int i) {
    const int4 r = s.rowinfo[i];
    PoaRow<SC> v;
    v.b = r.x; v.e = r.y; v.w = r.y - r.x + 1;
    v.c = reinterpret_cast<const SC *>(s.cells) + (((long long)r.w << 32) | (long long)(unsigned)r.z) * NPL;
    return v;
}
template <typename SC>
PD int poa_rcell(const PoaRow<SC> &v, int plane, int j, int inf_min) {
    if (j < v.b || j > v.e) return inf_min;
    return (int)v.c[(long long)plane * v.w + (j - v.b)];
}

// F1 / F2 of row i for the columns beg .. jmax, recomputed from the H / E1 / E2 planes of its predecessors exactly as the row DP
// computed them (pass 1: Hme = max(M, E1, E2) per column; pass 2: the prefix maxima, chunk by chunk with the same carries), as
// the DP would have STORED them ((SC) truncation included), into fc1 / fc2 (column - beg).  All lanes of the first warp.
// Why: the backtrack reads F only where it steps through an insertion (a few per cent of its steps), but keeping F1 / F2 of
// every cell is 2 of 5 planes of the scratch that limits how many heavy read sets run side by side.
template <typename SC>
PD void poa_fill_f(const PoaState &s, int i, int jmax, const uint8_t *q, int inf_min, int lane, int *fc1, int *fc2) {
    const int4 meta = s.rowmeta[i];
    const int nb = meta.y & 0xff, pn = meta.y >> 8, po = meta.z;
    const int4 ri = s.rowinfo[i];
    const int beg = ri.x, end = imin(ri.y, jmax);
    int carry_h = inf_min, carry_f1 = inf_min - GAP_OE1, carry_f2 = inf_min - GAP_OE2;
    for (int cb = beg; cb <= end; cb += POA_NL) {
        const int j = cb + lane;
        const bool valid = j <= end;
        int m = inf_min, e1in = inf_min, e2in = inf_min, sc = 0;
        if (valid) {
            for (int k = 0; k < pn; ++k) {
                const PoaRow<SC> pv = poa_row<SC>(s, s.edge_idx[po + k]);
                if (j - 1 >= pv.b && j - 1 <= pv.e) m = imax(m, (int)pv.c[j - 1 - pv.b]);
                if (j >= pv.b && j <= pv.e) { e1in = imax(e1in, (int)pv.c[pv.w + j - pv.b]); e2in = imax(e2in, (int)pv.c[2 * pv.w + j - pv.b]); }
            }
            if (j > 0) sc = poa_score(nb, q[j - 1]);
        }
        const int hme = valid ? imax(imax(m + sc, e1in), e2in) : inf_min;
        int left = poa_shfl_up(hme, 1);
        if (lane == 0) left = carry_h;
        int g1 = left - GAP_OE1 + lane * GAP_E1, g2 = left - GAP_OE2 + lane * GAP_E2;
#if POA_NL > 1
#pragma unroll
        for (int d = 1; d < POA_NL; d <<= 1) {
            const int t1 = poa_shfl_up(g1, d), t2 = poa_shfl_up(g2, d);
            if (lane >= d) { g1 = imax(g1, t1); g2 = imax(g2, t2); }
        }
#endif
        const int f1 = imax(g1 - lane * GAP_E1, carry_f1 - (lane + 1) * GAP_E1);
        const int f2 = imax(g2 - lane * GAP_E2, carry_f2 - (lane + 1) * GAP_E2);
        if (valid) { fc1[j - beg] = (int)(SC)f1; fc2[j - beg] = (int)(SC)f2; }
        carry_h = poa_shfl(hme, POA_NL - 1);
        carry_f1 = poa_shfl(f1, POA_NL - 1);
        carry_f2 = poa_shfl(f2, POA_NL - 1);
    }
}

// simd_abpoa_cg_backtrack (abpoa_align_simd.c:309-458) with put_gap_on_right = put_gap_at_end = 0.
// One step is a chain of dependent loads; it is kept short by reading what the topological sort and the DP left per INDEX
// (rowmeta: node id, base, in-edges; edge_idx: the predecessors' indices; rowinfo: band and cells) instead of going through
// the node arrays (idx2id -> in_off -> edge_id -> id2idx -> dp_beg / dp_end / row_off -> cell).
// All lanes of the first warp enter; lane 0 walks.  When a step reaches its F tests in a row whose F values are not at hand it
// stops, the warp recomputes them (poa_fill_f into the BFS scratch arrays deg / queue, free during an alignment) and lane 0
// takes the step again from its start -- the M and E tests it repeats have no side effects when they fail.
template <typename SC>
PD void poa_backtrack(PoaState &s, const uint8_t *q, int qlen, int inf_min, int lane) {
    int i = s.best_i, j = s.best_j, cur_op = OP_ALL;
    int guard = 2 * s.node_n + qlen + 8;
    int4 meta = s.rowmeta[i];
    PoaRow<SC> cur = poa_row<SC>(s, i);
    int *fc1 = s.deg, *fc2 = s.queue;
    int fc_row = -1;                                           // row whose F1 / F2 are in fc1 / fc2 (uniform)
    if (lane == 0) {
        s.n_cigar = 0;
        if (j < qlen) poa_push_cigar(s, CG_INS, qlen - j);
    }
    for (;;) {
        int need = 2;                                          // lane 0 decides: 1 = F values of row i up to column j wanted, 2 = done
        if (lane == 0)
        while (i > 0 && j > 0) {
            if (--guard < 0 || s.status) { if (!s.status) s.status = POA_ERR_BACKTRACK; break; }
            const int id = meta.x;
            const int sc = poa_score(meta.y & 0xff, q[j - 1]);
            const int pn = meta.y >> 8, po = meta.z;
            const int Hij = poa_rcell<SC>(cur, 0, j, inf_min);
            bool hit = false;
            if (cur_op & OP_M) {
                for (int k = 0; k < pn; ++k) {
                    const int p = s.edge_idx[po + k];
                    const PoaRow<SC> pv = poa_row<SC>(s, p);
                    const int4 pm = s.rowmeta[p];
                    if (j - 1 < pv.b || j - 1 > pv.e) continue;
                    if (poa_rcell<SC>(pv, 0, j - 1, inf_min) + sc == Hij) {
                        poa_push_cigar(s, CG_MATCH, id);
                        i = p; cur = pv; meta = pm; --j; hit = true; cur_op = OP_ALL;
                        break;
                    }
                }
            }
            if (!hit && (cur_op & OP_E)) {
                for (int k = 0; k < pn && !hit; ++k) {
                    const int p = s.edge_idx[po + k];
                    const PoaRow<SC> pv = poa_row<SC>(s, p);
                    const int4 pm = s.rowmeta[p];
                    if (j < pv.b || j > pv.e) continue;
                    const int pH = poa_rcell<SC>(pv, 0, j, inf_min);
                    if (cur_op & OP_E1) {
                        const int pe1 = poa_rcell<SC>(pv, 1, j, inf_min);
                        const bool ok = (cur_op & OP_M) ? (Hij == pe1) : (poa_rcell<SC>(cur, 1, j, inf_min) == pe1 - GAP_E1);
                        if (ok) {
                            cur_op = (pH - GAP_OE1 == pe1) ? (OP_M | OP_F) : OP_E1;
                            hit = true; poa_push_cigar(s, CG_DEL, id); i = p; cur = pv; meta = pm;
                            break;
                        }
                    }
                    if (cur_op & OP_E2) {
                        const int pe2 = poa_rcell<SC>(pv, 2, j, inf_min);
                        const bool ok = (cur_op & OP_M) ? (Hij == pe2) : (poa_rcell<SC>(cur, 2, j, inf_min) == pe2 - GAP_E2);
                        if (ok) {
                            cur_op = (pH - GAP_OE2 == pe2) ? (OP_M | OP_F) : OP_E2;
                            hit = true; poa_push_cigar(s, CG_DEL, id); i = p; cur = pv; meta = pm;
                            break;
                        }
                    }
                }
            }
            if (!hit && (cur_op & OP_F)) {
                if (fc_row != i) { ++guard; need = 1; break; }  // recompute this row's F1 / F2 up to column j, then take the step again
                // F1 / F2 of (i, j) and (i, j - 1): outside the row's band they read as inf_min, like every cell
                const bool in_j = j >= cur.b && j <= cur.e, in_l = j - 1 >= cur.b && j - 1 <= cur.e;
                const int Hl = poa_rcell<SC>(cur, 0, j - 1, inf_min);
                if (cur_op & OP_F1) {
                    const int f1 = in_j ? fc1[j - cur.b] : inf_min;
                    if (!(cur_op & OP_M) || Hij == f1) {
                        if (Hl - GAP_OE1 == f1) { cur_op = OP_M | OP_E; hit = true; }
                        else if ((in_l ? fc1[j - 1 - cur.b] : inf_min) - GAP_E1 == f1) { cur_op = OP_F1; hit = true; }
                    }
                }
                if (!hit && (cur_op & OP_F2)) {
                    const int f2 = in_j ? fc2[j - cur.b] : inf_min;
                    if (!(cur_op & OP_M) || Hij == f2) {
                        if (Hl - GAP_OE2 == f2) { cur_op = OP_M | OP_E; hit = true; }
                        else if ((in_l ? fc2[j - 1 - cur.b] : inf_min) - GAP_E2 == f2) { cur_op = OP_F2; hit = true; }
                    }
                }
                if (hit) { poa_push_cigar(s, CG_INS, 1); --j; }
            }
            if (!hit && (cur_op & OP_M)) {
                for (int k = 0; k < pn; ++k) {
                    const int p = s.edge_idx[po + k];
                    const PoaRow<SC> pv = poa_row<SC>(s, p);
                    const int4 pm = s.rowmeta[p];
                    if (j - 1 < pv.b || j - 1 > pv.e) continue;
                    if (poa_rcell<SC>(pv, 0, j - 1, inf_min) + sc == Hij) {
                        poa_push_cigar(s, CG_MATCH, id);
                        i = p; cur = pv; meta = pm; --j; hit = true; cur_op = OP_ALL;
                        break;
                    }
                }
            }
            if (!hit) { s.status = POA_ERR_BACKTRACK; break; }
        }
        need = poa_shfl(need, 0);
        if (need != 1) break;
        const int ri = poa_shfl(i, 0), rj = poa_shfl(j, 0);
        {   // fc1 / fc2 are node arrays: a band wider than the node capacity (only with a block forced too small) asks for a larger block
            const int4 rr = s.rowinfo[ri];
            if (imin(rr.y, rj) - rr.x + 1 > s.node_cap) { if (lane == 0) s.status = POA_ERR_NODES; break; }
        }
        poa_fill_f<SC>(s, ri, rj, q, inf_min, lane, fc1, fc2);
        fc_row = ri;
        poa_sync();
    }
    if (lane == 0 && !s.status && j > 0) poa_push_cigar(s, CG_INS, j);
}

// One alignment of q against the sorted graph: first row, row DP in index order, best end cell.  All lanes.
// Returns the number of band cells computed (uniform).  abpoa_align_simd.c:1200-1231
template <typename SC>
PD long long poa_align_dp(PoaState &s, const uint8_t *q, int qlen, int inf_min) {
    const int lane = poa_lane();
    const PoaState g = s;                                      // register copy of the pointer block (s lives in shared memory)
    const int n = g.node_n;
    const int w = EXTRA_B + (int)(EXTRA_F * (float)qlen);      // abpoa_align_simd.c:474
    SC *cells = reinterpret_cast<SC *>(g.cells);
    const long long cell_cap = g.cell_bytes / (long long)(NPL * sizeof(SC));
    long long top = 0;

    // ---- first row (simd_abpoa_cg_first_row / _first_dp, :669-688, :617-633)
    if (lane == 0) {
        g.max_left[SRC] = g.max_right[SRC] = 0;
        for (int i = 0; i < g.out_n[SRC]; ++i) { const int o = g.edge_id[g.out_off[SRC] + i]; g.max_left[o] = g.max_right[o] = 1; }
    }
    poa_sync();
    const int end0 = imin(qlen, imax(g.max_right[SRC], qlen - g.remain[SRC]) + w);
    {
        const int w0 = end0 + 1;
        if (w0 > cell_cap) { if (lane == 0) s.status = POA_ERR_CELLS; poa_sync(); return 0; }
        if (lane == 0) { g.dp_beg[0] = 0; g.dp_end[0] = end0; g.row_off[0] = 0; g.rowinfo[0] = make_int4(0, end0, 0, 0); }
        for (int j = lane; j <= end0; j += POA_NL) {
            int h, e1 = inf_min, e2 = inf_min, f1, f2;
            if (j == 0) { h = 0; e1 = -GAP_OE1; e2 = -GAP_OE2; f1 = f2 = inf_min; }
            else { f1 = -GAP_O1 - GAP_E1 * j; f2 = -GAP_O2 - GAP_E2 * j; h = imax(f1, f2); }
            cells[j] = (SC)h; cells[w0 + j] = (SC)e1; cells[2 * w0 + j] = (SC)e2; (void)f1; (void)f2;
        }
        top = w0;
    }
    poa_sync();

    // ---- rows 1 .. n-2 in topological index order (simd_abpoa_cg_dp, :935-1074).
    // Latency plan: a row's static description (rowmeta, predecessor indices) is fetched one and two rows ahead; the
    // previous row's band sits in registers (it is the predecessor of most rows); and a row is computed in blocks of
    // ROW_CB x 32 columns whose predecessor cells are all requested before the first scan starts, so a block pays one
    // memory round trip, not one per 32 columns.
    int lb = 0, le = end0;                                     // band of row i-1
    long long lo = 0;                                          // its first cell
    // The band hints (max_pos_left / max_pos_right of a node) are written by a row's predecessors and read when the row
    // starts; through memory that is a store -> load round trip at the head of every row and a chain of three dependent loads
    // (out-degree -> edge -> the target's hint) at its tail.  Instead: the next row's hints are requested one row ahead, as every
    // predecessor but the current row has left them, and the current row's own share is added in registers; the current
    // row's out-edge targets (rowout, lane k = edge k) and their hints are requested when the row starts.
    int4 meta = g.rowmeta[1], meta_n = g.rowmeta[imin(2, n - 1)];
    int4 om = g.rowout[1];
    int ml = g.max_left[meta.x], mr = g.max_right[meta.x];
    int pi[PRED_REG];
#pragma unroll
    for (int k = 0; k < PRED_REG; ++k) pi[k] = k < (meta.y >> 8) ? g.edge_idx[meta.z + k] : -1;
    for (int i = 1; i < n - 1; ++i) {
        const int id = meta.x, nb = meta.y & 0xff, pn = meta.y >> 8, po = meta.z;
        const int rt = qlen - meta.w;                          // remain[SINK] == -1 (abpoa_align.h:34-35)
        const int4 meta_nn = g.rowmeta[imin(i + 2, n - 1)];    // in flight during this row
        const int4 om_n = g.rowout[imin(i + 1, n - 1)];
        int ml_n = g.max_left[meta_n.x], mr_n = g.max_right[meta_n.x];
        int pi_n[PRED_REG];
#pragma unroll
        for (int k = 0; k < PRED_REG; ++k) pi_n[k] = k < (meta_n.y >> 8) ? g.edge_idx[meta_n.z + k] : -1;
        int to = -1, tml = 0, tmr = 0;                         // this lane's out-edge target and its hints so far
        if (lane < om.x) {
            to = lane == 0 ? om.z : (lane == 1 ? om.w : g.edge_id[om.y + lane]);
            tml = g.max_left[to]; tmr = g.max_right[to];
        }
        // the first PRED_REG predecessors' bands and row pointers stay in registers for the whole row (an empty slot
        // has an empty band, so it never contributes); nodes with more in-edges take the generic loop for the rest
        int pbv[PRED_REG], pev[PRED_REG];
        const SC *prv[PRED_REG];
#pragma unroll
        for (int k = 0; k < PRED_REG; ++k) {
            pbv[k] = 0x7fffffff; pev[k] = -2; prv[k] = cells;
            const int p = pi[k];
            if (p == i - 1) { pbv[k] = lb; pev[k] = le; prv[k] = cells + lo * NPL; }
            else if (p >= 0) { const int4 ri = g.rowinfo[p]; pbv[k] = ri.x; pev[k] = ri.y; prv[k] = cells + (((long long)ri.w << 32) | (long long)(unsigned)ri.z) * NPL; }
        }
        int beg = imax(0, imin(ml, rt) - w);
        const int end = imin(qlen, imax(mr, rt) + w);
        int min_pre_beg = 0x7fffffff;
#pragma unroll
        for (int k = 0; k < PRED_REG; ++k) min_pre_beg = imin(min_pre_beg, pbv[k]);
        for (int k = PRED_REG; k < pn; ++k) min_pre_beg = imin(min_pre_beg, g.dp_beg[g.edge_idx[po + k]]);
        if (beg < min_pre_beg) beg = min_pre_beg;              // :957-959 at vector width 1
        const int wd = end - beg + 1;
        if (wd <= 0 || pn <= 0) { if (lane == 0) s.status = POA_ERR_BAND; poa_sync(); return top; }
        if (top + wd > cell_cap) { if (lane == 0) s.status = POA_ERR_CELLS; poa_sync(); return top; }
        SC *row = cells + top * NPL;
        if (lane == 0) { g.dp_beg[i] = beg; g.dp_end[i] = end; g.row_off[i] = top; g.rowinfo[i] = make_int4(beg, end, (int)(unsigned)(top & 0xffffffffll), (int)(top >> 32)); }

        int carry_h = inf_min;                                 // Hme of the column left of the chunk
        int carry_f1 = inf_min - GAP_OE1, carry_f2 = inf_min - GAP_OE2;
        int bv = -2147483647 - 1, bl = -1, br = -1;            // this lane's row maximum, first and last column reaching it
        for (int cb = beg; cb <= end; cb += ROW_CB * POA_NL) {
            int hmev[ROW_CB], e1v[ROW_CB], e2v[ROW_CB];
            // pass 1: every predecessor cell of the block is requested here (independent loads)
#pragma unroll
            for (int u = 0; u < ROW_CB; ++u) {
                const int j = cb + u * POA_NL + lane;
                const bool valid = j <= end;
                int m = inf_min, e1in = inf_min, e2in = inf_min, sc = 0;
                if (valid) {
#pragma unroll
                    for (int k = 0; k < PRED_REG; ++k) {       // max over predecessors is order independent; empty slots stay predicated
                        const int pb = pbv[k], pe = pev[k], pw = pe - pb + 1;   // (a uniform branch around them: +8 % DP time, the loads stop overlapping)
                        const SC *pr = prv[k];
                        if (j - 1 >= pb && j - 1 <= pe) m = imax(m, (int)pr[j - 1 - pb]);
                        if (j >= pb && j <= pe) { e1in = imax(e1in, (int)pr[pw + j - pb]); e2in = imax(e2in, (int)pr[2 * pw + j - pb]); }
                    }
                    for (int k = PRED_REG; k < pn; ++k) {
                        const int p = g.edge_idx[po + k];
                        const int pb = g.dp_beg[p], pe = g.dp_end[p], pw = pe - pb + 1;
                        const SC *pr = cells + g.row_off[p] * NPL;
                        if (j - 1 >= pb && j - 1 <= pe) m = imax(m, (int)pr[j - 1 - pb]);
                        if (j >= pb && j <= pe) { e1in = imax(e1in, (int)pr[pw + j - pb]); e2in = imax(e2in, (int)pr[2 * pw + j - pb]); }
                    }
                    if (j > 0) sc = poa_score(nb, q[j - 1]);   // query profile column 0 is 0 (:520)
                }
                hmev[u] = valid ? imax(imax(m + sc, e1in), e2in) : inf_min;
                e1v[u] = e1in; e2v[u] = e2in;
            }
            // pass 2: F[j] = max(Hme[j-1] - oe, F[j-1] - e) as a prefix max of (Hme[k-1] - oe + k*e) across the lanes (:1052-1059).
            // 2a: the prefix maxima inside each 32-column chunk -- they need only Hme, so the chunks' shuffle chains are
            // independent and overlap; 2b: the running F of the columns to the left joins chunk by chunk (one shuffle each).
            int g1v[ROW_CB], g2v[ROW_CB];
#pragma unroll
            for (int u = 0; u < ROW_CB; ++u) {
                const int prev_last = u == 0 ? carry_h : poa_shfl(hmev[u > 0 ? u - 1 : 0], POA_NL - 1);   // all lanes shuffle
                int left = poa_shfl_up(hmev[u], 1);
                if (lane == 0) left = prev_last;
                int g1 = left - GAP_OE1 + lane * GAP_E1, g2 = left - GAP_OE2 + lane * GAP_E2;
#if POA_NL > 1
#pragma unroll
                for (int d = 1; d < POA_NL; d <<= 1) {
                    const int t1 = poa_shfl_up(g1, d), t2 = poa_shfl_up(g2, d);
                    if (lane >= d) { g1 = imax(g1, t1); g2 = imax(g2, t2); }
                }
#endif
                g1v[u] = g1; g2v[u] = g2;
            }
#pragma unroll
            for (int u = 0; u < ROW_CB; ++u) {
                const int c0 = cb + u * POA_NL;
                if (c0 > end) break;
                const int j = c0 + lane;
                const bool valid = j <= end;
                const int hme = hmev[u];
                const int f1 = imax(g1v[u] - lane * GAP_E1, carry_f1 - (lane + 1) * GAP_E1);
                const int f2 = imax(g2v[u] - lane * GAP_E2, carry_f2 - (lane + 1) * GAP_E2);
                const int h = imax(hme, imax(f1, f2));
                if (valid) {
                    const int x = j - beg;
                    row[x] = (SC)h;
                    row[wd + x] = (SC)imax(e1v[u] - GAP_E1, h - GAP_OE1);      // E for the next row (:1067-1071)
                    row[2 * wd + x] = (SC)imax(e2v[u] - GAP_E2, h - GAP_OE2);
                    if (h > bv) { bv = h; bl = br = j; } else if (h == bv) br = j;
                }
                carry_h = poa_shfl(hme, POA_NL - 1);
                carry_f1 = poa_shfl(f1, POA_NL - 1);
                carry_f2 = poa_shfl(f2, POA_NL - 1);
            }
        }
        // row arg-max -> band hints of the successors (simd_abpoa_max_in_row :1107-1119, simd_abpoa_ada_max_i :1121-1130)
        const int vmax = poa_max_all(bv);
        int left_i = -1, right_i = -1;
        if (vmax >= inf_min) {
            right_i = poa_max_all(bv == vmax ? br : -1);
            if (vmax > inf_min) left_i = poa_min_all(bv == vmax ? bl : 0x7fffffff);
        }
        bool feeds_next = to == meta_n.x;                      // (to == -1 never matches a node id)
        if (to >= 0) {
            if (right_i + 1 > tmr) g.max_right[to] = right_i + 1;
            if (left_i + 1 < tml) g.max_left[to] = left_i + 1;
        }
        for (int k = lane + POA_NL; k < om.x; k += POA_NL) {   // more out-edges than lanes
            const int o = g.edge_id[om.y + k];
            if (right_i + 1 > g.max_right[o]) g.max_right[o] = right_i + 1;
            if (left_i + 1 < g.max_left[o]) g.max_left[o] = left_i + 1;
            feeds_next |= o == meta_n.x;
        }
        if (poa_any(feeds_next)) { mr_n = imax(mr_n, right_i + 1); ml_n = imin(ml_n, left_i + 1); }
        ml = ml_n; mr = mr_n; om = om_n;
        lb = beg; le = end; lo = top;
        top += wd;
        meta = meta_n; meta_n = meta_nn;
#pragma unroll
        for (int k = 0; k < PRED_REG; ++k) pi[k] = pi_n[k];
        poa_sync();
    }
    // ---- best end cell among SINK's predecessors (simd_abpoa_global_get_max, :1092-1105)
    if (lane == 0) {
        int best = inf_min, bi = 0, bj = 0;
        for (int k = 0; k < g.in_n[SINK]; ++k) {
            const int p = g.id2idx[g.edge_id[g.in_off[SINK] + k]];
            const int e = qlen > g.dp_end[p] ? g.dp_end[p] : qlen;
            const int v = poa_cell<SC>(g, p, 0, e, inf_min);
            if (v > best) { best = v; bi = p; bj = e; }
        }
        s.best_score = best; s.best_i = bi; s.best_j = bj;
    }
    poa_sync();
    return top;
}

// cross-warp exchange of one CTA-wide row (shared memory): lane 31's Hme and scan totals, per-warp row maxima
struct PoaXchg {
    int h[8], g1[8], g2[8];
    int bv[8], bl[8], br[8];
};

#ifndef POA_HOSTSIM
// CTA-wide row DP for the large read sets: NW warps share a row, thread t owns column cb + t.  Same recurrences as
// poa_align_dp; F1/F2 are written as one prefix max over the whole row of g(k) = Hme[k-1] - oe + k*e
// (F[j] = max_{k<=j} g(k) - j*e, identical in exact integer arithmetic to F[j] = max(Hme[j-1] - oe, F[j-1] - e)):
// warp shuffle scan, lane-31 totals through shared memory, running maximum carried across 32*NW-column blocks.
template <typename SC, int NW>
__device__ __noinline__ long long poa_align_dp_cta(PoaState &s, PoaXchg &x, const uint8_t *q, int qlen, int inf_min) {
    constexpr int NT = 32 * NW;
    constexpr int IMIN = -2147483647 - 1;
    const int t = threadIdx.x, lane = t & 31, wp = t >> 5;
    const PoaState g = s;
    const int n = g.node_n;
    const int w = EXTRA_B + (int)(EXTRA_F * (float)qlen);
    SC *cells = reinterpret_cast<SC *>(g.cells);
    const long long cell_cap = g.cell_bytes / (long long)(NPL * sizeof(SC));
    long long top = 0;

    if (t == 0) {
        g.max_left[SRC] = g.max_right[SRC] = 0;
        for (int i = 0; i < g.out_n[SRC]; ++i) { const int o = g.edge_id[g.out_off[SRC] + i]; g.max_left[o] = g.max_right[o] = 1; }
    }
    __syncthreads();
    const int end0 = imin(qlen, imax(g.max_right[SRC], qlen - g.remain[SRC]) + w);
    {
        const int w0 = end0 + 1;
        if (w0 > cell_cap) { if (t == 0) s.status = POA_ERR_CELLS; __syncthreads(); return 0; }
        if (t == 0) { g.dp_beg[0] = 0; g.dp_end[0] = end0; g.row_off[0] = 0; g.rowinfo[0] = make_int4(0, end0, 0, 0); }
        for (int j = t; j <= end0; j += NT) {
            int h, e1 = inf_min, e2 = inf_min, f1, f2;
            if (j == 0) { h = 0; e1 = -GAP_OE1; e2 = -GAP_OE2; f1 = f2 = inf_min; }
            else { f1 = -GAP_O1 - GAP_E1 * j; f2 = -GAP_O2 - GAP_E2 * j; h = imax(f1, f2); }
            cells[j] = (SC)h; cells[w0 + j] = (SC)e1; cells[2 * w0 + j] = (SC)e2; (void)f1; (void)f2;
        }
        top = w0;
    }
    __syncthreads();

    int lb = 0, le = end0;
    long long lo = 0;
    int4 meta = g.rowmeta[1], meta_n = g.rowmeta[imin(2, n - 1)];
    int pi[PRED_REG];
#pragma unroll
    for (int k = 0; k < PRED_REG; ++k) pi[k] = k < (meta.y >> 8) ? g.edge_idx[meta.z + k] : -1;
    for (int i = 1; i < n - 1; ++i) {
        const int id = meta.x, nb = meta.y & 0xff, pn = meta.y >> 8, po = meta.z;
        const int rt = qlen - meta.w;
        const int4 meta_nn = g.rowmeta[imin(i + 2, n - 1)];
        int pi_n[PRED_REG];
#pragma unroll
        for (int k = 0; k < PRED_REG; ++k) pi_n[k] = k < (meta_n.y >> 8) ? g.edge_idx[meta_n.z + k] : -1;
        const int ml = g.max_left[id], mr = g.max_right[id];
        int pbv[PRED_REG], pev[PRED_REG];
        const SC *prv[PRED_REG];
#pragma unroll
        for (int k = 0; k < PRED_REG; ++k) {
            pbv[k] = 0x7fffffff; pev[k] = -2; prv[k] = cells;
            const int p = pi[k];
            if (p == i - 1) { pbv[k] = lb; pev[k] = le; prv[k] = cells + lo * NPL; }
            else if (p >= 0) { const int4 ri = g.rowinfo[p]; pbv[k] = ri.x; pev[k] = ri.y; prv[k] = cells + (((long long)ri.w << 32) | (long long)(unsigned)ri.z) * NPL; }
        }
        int beg = imax(0, imin(ml, rt) - w);
        const int end = imin(qlen, imax(mr, rt) + w);
        int min_pre_beg = 0x7fffffff;
#pragma unroll
        for (int k = 0; k < PRED_REG; ++k) min_pre_beg = imin(min_pre_beg, pbv[k]);
        for (int k = PRED_REG; k < pn; ++k) min_pre_beg = imin(min_pre_beg, g.dp_beg[g.edge_idx[po + k]]);
        if (beg < min_pre_beg) beg = min_pre_beg;
        const int wd = end - beg + 1;
        if (wd <= 0 || pn <= 0) { if (t == 0) s.status = POA_ERR_BAND; __syncthreads(); return top; }
        if (top + wd > cell_cap) { if (t == 0) s.status = POA_ERR_CELLS; __syncthreads(); return top; }
        SC *row = cells + top * NPL;
        if (t == 0) { g.dp_beg[i] = beg; g.dp_end[i] = end; g.row_off[i] = top; g.rowinfo[i] = make_int4(beg, end, (int)(unsigned)(top & 0xffffffffll), (int)(top >> 32)); }

        int carry_h = inf_min;                                 // Hme of the column left of the block
        int run1 = IMIN, run2 = IMIN;                          // max of g1 / g2 over all columns left of the block
        int bv = IMIN, bl = -1, br = -1;
        for (int cb = beg; cb <= end; cb += NT) {
            const int j = cb + t;
            const bool valid = j <= end;
            int m = inf_min, e1in = inf_min, e2in = inf_min, sc = 0;
            if (valid) {
#pragma unroll
                for (int k = 0; k < PRED_REG; ++k) {
                    const int pb = pbv[k], pe = pev[k], pw = pe - pb + 1;
                    const SC *pr = prv[k];
                    if (j - 1 >= pb && j - 1 <= pe) m = imax(m, (int)pr[j - 1 - pb]);
                    if (j >= pb && j <= pe) { e1in = imax(e1in, (int)pr[pw + j - pb]); e2in = imax(e2in, (int)pr[2 * pw + j - pb]); }
                }
                for (int k = PRED_REG; k < pn; ++k) {
                    const int p = g.edge_idx[po + k];
                    const int pb = g.dp_beg[p], pe = g.dp_end[p], pw = pe - pb + 1;
                    const SC *pr = cells + g.row_off[p] * NPL;
                    if (j - 1 >= pb && j - 1 <= pe) m = imax(m, (int)pr[j - 1 - pb]);
                    if (j >= pb && j <= pe) { e1in = imax(e1in, (int)pr[pw + j - pb]); e2in = imax(e2in, (int)pr[2 * pw + j - pb]); }
                }
                if (j > 0) sc = poa_score(nb, q[j - 1]);
            }
            const int hme = valid ? imax(imax(m + sc, e1in), e2in) : inf_min;
            int left = __shfl_up_sync(0xffffffffu, hme, 1);
            if (lane == 31) x.h[wp] = hme;
            __syncthreads();                                   // (A) lane-31 Hme of every warp visible
            if (lane == 0) left = wp == 0 ? carry_h : x.h[wp - 1];
            carry_h = x.h[NW - 1];
            int g1 = left - GAP_OE1 + j * GAP_E1, g2 = left - GAP_OE2 + j * GAP_E2;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const int t1 = __shfl_up_sync(0xffffffffu, g1, d), t2 = __shfl_up_sync(0xffffffffu, g2, d);
                if (lane >= d) { g1 = imax(g1, t1); g2 = imax(g2, t2); }
            }
            if (lane == 31) { x.g1[wp] = g1; x.g2[wp] = g2; }
            __syncthreads();                                   // (B) warp totals visible
            int pre1 = run1, pre2 = run2;
#pragma unroll
            for (int v = 0; v < NW; ++v) {
                const int a1 = x.g1[v], a2 = x.g2[v];
                if (v < wp) { pre1 = imax(pre1, a1); pre2 = imax(pre2, a2); }
                run1 = imax(run1, a1); run2 = imax(run2, a2);
            }
            const int f1 = imax(g1, pre1) - j * GAP_E1;
            const int f2 = imax(g2, pre2) - j * GAP_E2;
            const int h = imax(hme, imax(f1, f2));
            if (valid) {
                const int xo = j - beg;
                row[xo] = (SC)h;
                row[wd + xo] = (SC)imax(e1in - GAP_E1, h - GAP_OE1);
                row[2 * wd + xo] = (SC)imax(e2in - GAP_E2, h - GAP_OE2);
                if (h > bv) { bv = h; bl = br = j; } else if (h == bv) br = j;
            }
        }
        // row arg-max across the CTA
        {
            const int wv = __reduce_max_sync(0xffffffffu, bv);
            const int wl = __reduce_min_sync(0xffffffffu, bv == wv ? bl : 0x7fffffff);
            const int wr = __reduce_max_sync(0xffffffffu, bv == wv ? br : -1);
            if (lane == 0) { x.bv[wp] = wv; x.bl[wp] = wl; x.br[wp] = wr; }
        }
        __syncthreads();                                       // (C)
        int vmax = IMIN;
#pragma unroll
        for (int v = 0; v < NW; ++v) vmax = imax(vmax, x.bv[v]);
        int left_i = 0x7fffffff, right_i = -1;
#pragma unroll
        for (int v = 0; v < NW; ++v) if (x.bv[v] == vmax) { left_i = imin(left_i, x.bl[v]); right_i = imax(right_i, x.br[v]); }
        if (vmax < inf_min) { left_i = -1; right_i = -1; }
        else if (vmax == inf_min || left_i == 0x7fffffff) left_i = -1;
        for (int k = t; k < g.out_n[id]; k += NT) {
            const int o = g.edge_id[g.out_off[id] + k];
            if (right_i + 1 > g.max_right[o]) g.max_right[o] = right_i + 1;
            if (left_i + 1 < g.max_left[o]) g.max_left[o] = left_i + 1;
        }
        lb = beg; le = end; lo = top;
        top += wd;
        meta = meta_n; meta_n = meta_nn;
#pragma unroll
        for (int k = 0; k < PRED_REG; ++k) pi[k] = pi_n[k];
        __syncthreads();                                       // (D) row visible to the next one; exchange slots free
    }
    if (t == 0) {
        int best = inf_min, bi = 0, bj = 0;
        for (int k = 0; k < g.in_n[SINK]; ++k) {
            const int p = g.id2idx[g.edge_id[g.in_off[SINK] + k]];
            const int e = qlen > g.dp_end[p] ? g.dp_end[p] : qlen;
            const int v = poa_cell<SC>(g, p, 0, e, inf_min);
            if (v > best) { best = v; bi = p; bj = e; }
        }
        s.best_score = best; s.best_i = bi; s.best_j = bj;
    }
    __syncthreads();
    return top;
}
#endif

template <typename SC, int NW>
PD long long poa_align_any(PoaState &s, PoaXchg &x, const uint8_t *q, int qlen, int inf_min) {
#ifndef POA_HOSTSIM
    if (NW > 1) return poa_align_dp_cta<SC, NW>(s, x, q, qlen, inf_min);
#endif
    (void)x;
    return poa_align_dp<SC>(s, q, qlen, inf_min);
}

PD uint64_t poa_fnv(uint64_t h, int32_t v) {
    for (int k = 0; k < 4; ++k) { h ^= (uint64_t)(((uint32_t)v >> (8 * k)) & 0xffu); h *= 1099511628211ull; }
    return h;
}

// per-read observations for the parity tests (all optional)
struct PoaTrace {
    int32_t *best_score, *node_n, *n_ops;
    uint64_t *ops_hash;
    long long *cells;
    int32_t dump_read;          // global read index whose expanded ops go to dump (pairs op, node), or -1
    int32_t *dump;
    int32_t dump_cap;
};

// heaviest bundling (abpoa_heaviest_bundling, abpoa_output.c:477-547) + walk (abpoa_set_hb_cons :375-391).  lane 0.
// score -> max_left, max_out_id -> max_right (free after the last alignment).  Returns the consensus length.
PD int poa_consensus(PoaState &s, uint8_t *cons, int32_t *cov, int cap) {
    int *score = s.max_left, *max_out = s.max_right;
    for (int i = 0; i < s.node_n; ++i) s.deg[i] = s.out_n[i];
    int head = 0, tail = 0;
    s.queue[tail++] = SINK;
    bool done = false;
    while (head < tail && !done) {
        const int cur = s.queue[head++];
        if (cur == SINK) { max_out[cur] = -1; score[cur] = 0; }
        else if (cur == SRC) {
            int path_score = -1, path_max_w = -1, max_id = -1;
            for (int i = 0; i < s.out_n[cur]; ++i) {
                const int o = s.edge_id[s.out_off[cur] + i], ow = s.edge_w[s.out_off[cur] + i];
                if (ow > path_max_w || (ow == path_max_w && score[o] > path_score)) { max_id = o; path_score = score[o]; path_max_w = ow; }
            }
            max_out[cur] = max_id;
            done = true;
            break;
        } else {
            int max_w = -2147483647 - 1, max_id = -1;
            for (int i = 0; i < s.out_n[cur]; ++i) {
                const int o = s.edge_id[s.out_off[cur] + i], ow = s.edge_w[s.out_off[cur] + i];
                if (max_w < ow) { max_w = ow; max_id = o; }
                else if (max_w == ow && score[max_id] <= score[o]) max_id = o;
            }
            if (max_id < 0) { s.status = POA_ERR_CONS; return 0; }
            score[cur] = max_w + score[max_id];
            max_out[cur] = max_id;
        }
        for (int i = 0; i < s.in_n[cur]; ++i) {
            const int in = s.edge_id[s.in_off[cur] + i];
            if (--s.deg[in] == 0) s.queue[tail++] = in;
        }
    }
    if (!done) { s.status = POA_ERR_CONS; return 0; }
    int len = 0, cur = max_out[SRC];
    while (cur != SINK) {
        if (cur < 0 || len > s.node_n) { s.status = POA_ERR_CONS; return 0; }
        if (len < cap) { cons[len] = s.base[cur]; if (cov) cov[len] = s.n_read[cur]; }
        ++len;
        cur = max_out[cur];
    }
    return len;
}

// The whole read set: reads r0 .. r1 of the batch (bases[seq_off[r] .. seq_off[r+1])).  All lanes.  Returns the
// consensus length (uniform), 0 on error (s.status says why).
template <int NW, typename XCH>
PD int poa_run_set(PoaState &s, XCH &xch, long long r0, long long r1, const long long *seq_off, const uint8_t *bases, uint8_t *cons,
                   int32_t *cov, int cons_cap, const PoaTrace *tr) {
    const int lane = poa_tid<NW>();
    if (lane == 0) {                                           // abpoa_reset (abpoa_graph.c:782): graph = {SRC, SINK}
        s.node_n = 0; s.edge_top = 0; s.status = POA_OK; s.n_cigar = 0; s.cells_done = 0; s.cells_full = 0; s.cells_peak = 0;
        s.t_phase[0] = s.t_phase[1] = s.t_phase[2] = s.t_phase[3] = 0;
        poa_add_node(s, 0); poa_add_node(s, 0);
    }
    poa_bar<NW>();
    for (long long r = r0; r < r1 && !s.status; ++r) {
        const uint8_t *q = bases + seq_off[r];
        const int qlen = (int)(seq_off[r + 1] - seq_off[r]);
        if (qlen <= 0) { if (lane == 0) s.status = POA_ERR_INPUT; poa_bar<NW>(); break; }
        long long cells = 0;
        const int rows = s.node_n;
        if (rows <= 2) {                                       // abpoa_align_sequence_to_graph returns -1 (abpoa_align.c:201)
            if (lane == 0) {
                PoaState g = s;
                g.n_cigar = 0; g.best_score = 0;
                poa_add_first(g, q, qlen);
                s = g;
            }
        } else {
            // score width (abpoa_align_simd.c:1291-1302)
            const int len = qlen > rows ? qlen : rows;
            const int max_score = imax(qlen * MATCH, len * GAP_E1 + GAP_O1);
            const bool use16 = max_score <= 32767 - MISMATCH - GAP_OE1 - GAP_OE2;
            const long long t0 = poa_clock();
            if (use16) cells = poa_align_any<int16_t, NW>(s, xch, q, qlen, INF_MIN16);
            else cells = poa_align_any<int32_t, NW>(s, xch, q, qlen, INF_MIN32);
            const long long t1 = poa_clock();
            if (lane < POA_NL && !s.status) {                  // the first warp: lane 0 walks back, the others help with F1 / F2
                PoaState g = s;
                if (use16) poa_backtrack<int16_t>(g, q, qlen, INF_MIN16, lane); else poa_backtrack<int32_t>(g, q, qlen, INF_MIN32, lane);
                if (lane == 0) {
                    s.n_cigar = g.n_cigar; s.status = g.status;
                    s.t_phase[0] += t1 - t0; s.t_phase[1] += poa_clock() - t1;
                    s.cells_done += cells;
                    if (cells > s.cells_peak) s.cells_peak = cells;
                    s.cells_full += (long long)(rows - 1) * (long long)(qlen + 1);
                }
            }
        }
        poa_bar<NW>();
        if (lane == 0 && !s.status) {
            if (tr && tr->ops_hash) {                          // forward (query) order, as the reference's reversed cigar
                uint64_t h = 1469598103934665603ull;
                int nops = 0;
                for (int c = s.n_cigar - 1; c >= 0; --c) {
                    const int op = s.cigar[c].x, arg = s.cigar[c].y;
                    const int rep = op == CG_INS ? arg : 1;
                    for (int l = 0; l < rep; ++l) {
                        const int node = op == CG_INS ? -1 : arg;
                        h = poa_fnv(poa_fnv(h, op), node);
                        if (tr->dump && r == tr->dump_read && 2 * nops + 1 < tr->dump_cap) { tr->dump[2 * nops] = op; tr->dump[2 * nops + 1] = node; }
                        ++nops;
                    }
                }
                tr->ops_hash[r] = h; tr->n_ops[r] = nops;
            }
            if (tr && tr->best_score) { tr->best_score[r] = rows > 2 ? s.best_score : 0; tr->cells[r] = cells; }
        }
        if (lane < POA_NL && rows > 2 && !s.status) {          // the first warp: lane 0 fuses, the others prefetch for it
            const long long t2 = poa_clock();
            PoaState g = s;
            poa_add_alignment(g, q, qlen, lane);
            if (lane == 0) { s.node_n = g.node_n; s.edge_top = g.edge_top; s.status = g.status; s.t_phase[2] += poa_clock() - t2; }
        }
        poa_bar<NW>();
        if (s.status) break;
        const long long t3 = poa_clock();
        poa_topological_sort<NW>(s);
        if (lane == 0) s.t_phase[3] += poa_clock() - t3;
        if (lane == 0 && tr && tr->node_n) tr->node_n[r] = s.node_n;
        poa_bar<NW>();
    }
    int len = 0;
    if (lane == 0 && !s.status && s.node_n > 2) {
        PoaState g = s;
        len = poa_consensus(g, cons, cov, cons_cap);
        s.status = g.status;
    }
    if (lane == 0) s.ret_len = len;
    poa_bar<NW>();
    len = s.ret_len;
    poa_bar<NW>();
    return s.status ? 0 : len;
}

}  // namespace svtk_poa
