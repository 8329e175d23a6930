// svtk_launch.h -- host-side launch interface between the C ABI (svtk_capi.cpp) and the kernels (svtk_kernels.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace svtk {

// function attributes (dynamic shared memory limits) are per device: one process may drive several GPUs, one ctx each
inline bool first_use_on_device(unsigned long long &seen) {
    int d = 0;
    cudaGetDevice(&d);
    const unsigned long long bit = 1ull << (d & 63);
    if (seen & bit) return false;
    seen |= bit;
    return true;
}

struct LaunchReads {            // device pointers
    int n_tid;
    long long n_reads;
    const long long *tid_off;
    const uint32_t *pos, *end, *n_cigar;
    const unsigned long long *cigar_off;
    const uint32_t *cigar;
    uint32_t max_span;
    const uint32_t *tiles;          // position tiles built by launch_pos_tiles (nullptr: the kernels fall back to a search over pos)
    const long long *tile_off;
};

struct LaunchWindows {          // device pointers
    int n_win;
    const int32_t *chrom;
    const uint32_t *start, *end, *pos;
    const uint8_t *kind;
};

struct LaunchIndex {            // the signature index built by k_index_reads (device pointers)
    void *events;               // uint4 per event; index_pools() sub-pools of pool_cap events each, then index_fixed_slots() per chunk
    void *chunks;               // uint4 per chunk of index_reads_per_chunk() reads
    void *segs;                 // uint4 per extra segment
    uint32_t *rend;             // per read
    unsigned int *ctrl;         // index_ctrl_words() u32: [p * stride] sub-pool tops, [pools * stride] segment top, then error bits
    unsigned int pool_cap, seg_cap;
    int n_chunks;
    unsigned int *next_ctrl;    // control block of the NEXT build, cleared by this build's kernel (nullptr: the host clears it)
    int next_words;
};

int index_reads_per_chunk();
int index_pools();
int index_fixed_slots();      // event slots per chunk behind the sub-pools (0: none); their counts live in ctrl[p * stride + 1]
int index_ctrl_words();
int index_zeroes_next_ctrl();  // 1: launch_index_reads clears LaunchIndex::next_ctrl itself
int index_ctrl_stride();
int refine_block_segment_ints();
cudaError_t launch_index_reads(const LaunchReads &r, const LaunchIndex &x, cudaStream_t s);
cudaError_t launch_refine_indexed(const LaunchReads &r, const LaunchWindows &w, const LaunchIndex &x, int2 *range, int range_, int interval,
                                  int min_count, int32_t *result, int *ov_count, int2 *ov_items, int ov_capacity,
                                  unsigned long long *counters, cudaStream_t s);
cudaError_t launch_refine_block(const LaunchReads &r, const LaunchWindows &w, const LaunchIndex &x, const int2 *range, int range_,
                                int interval, int min_count, int32_t *result, int *ov_count, int2 *ov_items, int ov_capacity, int n_items,
                                int *arena, unsigned long long arena_ints, unsigned long long *arena_top, int *fail_flag, cudaStream_t s);
cudaError_t launch_check_layout(const LaunchReads &r, unsigned long long n_words, unsigned int *flag, cudaStream_t s);
int pos_tile_shift();
cudaError_t launch_last_pos(const LaunchReads &r, uint32_t *out, cudaStream_t s);      // out[t] = pos of reference t's last read (0xffffffff: none)
cudaError_t launch_pos_tiles(const LaunchReads &r, uint32_t *tiles, const long long *tile_off, cudaStream_t s);

int big_segment_ints();
int consensus_segment_smem_ints();

cudaError_t launch_max_span(const uint32_t *pos, const uint32_t *end, long long n, uint32_t *out, cudaStream_t s);
cudaError_t launch_plan(const LaunchReads &r, const LaunchWindows &w, int2 *range, cudaStream_t s);
cudaError_t launch_refine_fused(const LaunchReads &r, const LaunchWindows &w, const int2 *range, int range_, int interval,
                                int min_count, int32_t *result, int *ov_count, int2 *ov_items, int ov_capacity,
                                unsigned long long *counters, cudaStream_t s);
cudaError_t launch_refine_overflow(const LaunchReads &r, const LaunchWindows &w, const int2 *range, int range_, int interval,
                                   int min_count, int32_t *result, int *ov_count, int2 *ov_items, int ov_capacity, int n_items,
                                   int *arena, unsigned long long arena_ints, unsigned long long *arena_top, int *fail_flag,
                                   cudaStream_t s);
cudaError_t launch_consensus_segments(const int32_t *values, const long long *offsets, const int32_t *pos, int n_seg, int range_,
                                      int interval, int min_count, int *scratch, const long long *scratch_off, int32_t *result,
                                      cudaStream_t s);
cudaError_t launch_slide_tiles(const LaunchReads &r, const LaunchWindows &w, const int2 *range, int window_size, int slide,
                               int min_count, int *gseg, long long gseg_stride, int2 *out, int *need, cudaStream_t s);
cudaError_t launch_window_signatures(const LaunchReads &r, const LaunchWindows &w, const int2 *range, int *buf, int cap, int *n_out,
                                     cudaStream_t s);

}  // namespace svtk
