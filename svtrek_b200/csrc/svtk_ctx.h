// svtk_ctx.h -- the context behind svtk_ctx*: one GPU, one stream, the resident read set and reusable buffers.
#pragma once
#include "../../include/svtk.h"
#include "svtk_launch.h"

#include <cstddef>
#include <cstdint>
#include <string>
#include <vector>

namespace svtk {

int fail(int code, const char *fmt, ...);

struct DevBuf {
    void *p = nullptr;
    size_t cap = 0;
    int reserve(size_t bytes);   // grow-only; contents are NOT preserved
    void release();
};

struct PinBuf {
    void *p = nullptr;
    size_t cap = 0;
    int reserve(size_t bytes);
    void release();
};

struct Ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = true;
    svtk_params params{};

    // resident read set (HBM).  r_* own the memory when loaded from the host; attach leaves them empty.
    bool have_reads = false;
    LaunchReads reads{};
    std::vector<int64_t> h_tid_off;
    DevBuf r_tid_off, r_pos, r_end, r_ncig, r_off, r_cigar;
    DevBuf pos_tiles, pos_tile_off;       // position tiles of the resident reads (svtk_capi.cpp finish_reads)

    // signature index over the resident reads (svtk_index.cu): rebuilt by the first refine after a load
    long long n_words = 0;
    bool index_valid = false;
    bool index_fresh = false;      // built by the batch in flight: its error bits still have to be checked at sync
    LaunchIndex idx{};
    DevBuf idx_events, idx_chunks, idx_segs, idx_rend, idx_ctrl;      // idx_ctrl: index counters, then the batch flags (svtk_capi.cpp ensure_ctl)
    uint64_t idx_n_events = 0, idx_n_segs = 0;

    // per-batch scratch
    DevBuf range;        // int2 per window: candidate read range
    DevBuf ov_items;     // int2 per window: (window, signature count) of fast-path overflows
    DevBuf arena;        // global signature segments for windows beyond the big shared segment
    DevBuf counters;     // WorkCounters
    DevBuf scratch4;
    DevBuf win_blob, result;       // device copies for the host-buffer API
    DevBuf tmp_a, tmp_b, tmp_c;
    PinBuf stage[2];
    cudaEvent_t stage_ev[2] = {nullptr, nullptr};
    PinBuf h_idx, h_win, h_result;      // h_idx mirrors idx_ctrl: index counters, then the batch flags

    // a launched batch whose overflow windows have not been resolved yet
    bool pending = false;
    int pend_mode = 0;
    LaunchWindows pend_win{};
    int32_t *pend_result = nullptr;
    int last_overflow = 0;

    uint64_t launches = 0;
    uint64_t h2d_bytes = 0, d2h_bytes = 0;

    // svtk_audit_device: steps launched and not yet checked by svtk_sync; each owns a slot of h_audit (its control block copy)
    struct AuditStep { LaunchWindows win; int32_t *result; int slot; };
    std::vector<AuditStep> audit_q;
    PinBuf h_audit;

    // K3 (svtk_poa.cu): device copies of one batch of read sets, per-worker scratch, events
    DevBuf poa_in, poa_out, poa_scratch, poa_ctl;
    PinBuf poa_h;
    cudaEvent_t poa_ev[3] = {nullptr, nullptr, nullptr};
    cudaStream_t poa_side = nullptr;      // the CTA-wide launch of the heaviest read sets runs beside the warp-per-set launch

    // svtk_load_bam (svtk_bam.cu): the raw file, its inflated image and the record table stay resident (INS-sequence extraction
    // reads SEQ from them); reference names / lengths of the BAM header for the host front end
    DevBuf bam_comp, bam_infl, bam_blocks, bam_splits, bam_seg, bam_rec, bam_ctl;
    PinBuf bam_h;
    cudaEvent_t bam_ev[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
    // the loader's pipeline: pinned ring the file is read into, side streams the inflate launches run on, the pinned block table
    PinBuf bam_stage[3], bam_tab;
    DevBuf bam_tok, bam_ntok;             // match tokens between the two inflate phases, and their count per block
    cudaEvent_t bam_stage_ev[3] = {nullptr, nullptr, nullptr}, bam_side_ev[4] = {nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t bam_chunk_ev[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    cudaStream_t bam_side[4] = {nullptr, nullptr, nullptr, nullptr};
    std::vector<std::string> bam_ref_names;
    std::vector<uint32_t> bam_ref_lens;
    unsigned long long bam_infl_bytes = 0;
    long long bam_n_rec = 0;
    bool reads_from_bam = false;          // the resident reads are the records of the resident BAM image (index r == record r)
    // svtk_ins_sequences (svtk_ins.cu): the last batch of read sets, host side, owned by the ctx
    std::vector<int64_t> ins_set_off, ins_seq_off;
    std::vector<uint8_t> ins_bases;
    std::vector<int32_t> ins_window, ins_read;
    std::vector<uint32_t> ins_ref_pos;

    // live kernel timing (svtk_set_profiling): CUDA events around the fused refinement kernel on the ctx stream
    bool profile = false, prof_pending = false, prof_idx_pending = false;
    bool flags_clean = false;                  // the current block's batch flags are zero (a build cleared them, no batch ran since)
    int ctl_cur = 0;                           // which of the two control blocks the current index / batch uses
    bool ctl_clean[2] = {false, false};        // block is known to hold zeros (cleared by the index kernel of the previous build)
    int prof_stride = 1;                       // bracket every prof_stride-th index build / refine launch (svtk_set_profiling(ctx, n))
    unsigned long long prof_tick_idx = 0, prof_tick_ref = 0;
    cudaEvent_t prof_ev[4] = {nullptr, nullptr, nullptr, nullptr};     // [0,1] refine kernel, [2,3] index kernel
    double refine_ms = 0.0, index_ms = 0.0;
    uint64_t refine_launches = 0, index_launches = 0;
};

int upload(Ctx *c, void *dst, const void *src, size_t bytes);   // host -> device on the ctx stream (pinned staging)
int set_device(Ctx *c);
int adopt_device_reads(Ctx *c, const svtk_reads *d);
int bam_window_weights(const char *bam_path, const char *bai_path, int n, const int32_t *chrom, const uint32_t *start, const uint32_t *end, double *weight);   // svtk_bam.cu   // what svtk_attach_reads_device does, for loaders that fill c->r_* themselves

}  // namespace svtk
