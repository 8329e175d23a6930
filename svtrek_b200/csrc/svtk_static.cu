// svtk_static.cu -- svtk_consensus_pos_batch_static: consensus_pos on caller-supplied DEVICE arrays with no allocation at all.
//
// No counterpart in the reference: added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  It is the k_consensus_segments kernel of
// svtk_kernels.cu (parity aid for consensus_pos, refinement.c:41-101) with its global sort scratch -- the segments too
// long for the 8 192-int shared-memory segment -- moved from the ctx arena (cudaMalloc'ed, grown on demand) into a
// file-scope device array of fixed size.  Every GPU that loads the library holds that array (8 MiB) until exit.
//
// Scratch layout: segment s = values[b .. b + n) sorts in g_static_scratch[2b .. 2b + next_pow2(n)).  next_pow2(n) < 2n, so
// segments never overlap and no bump counter is needed; every word a segment reads back (its values and the sort's padding) is
// written by the same CTA in the same launch, so nothing has to start at zero and the previous call's contents do not
// matter.  All calls on one GPU share the array: two calls that overlap on the GPU (different streams, different ctx) race.
#include "svtk_ctx.h"
#include "svtk_device.cuh"

#include <mutex>

namespace svtk {

constexpr int STATIC_NT = 256;
constexpr int STATIC_CAP_SMEM = 8192;                                    // ints; same split as k_consensus_segments
constexpr long long STATIC_MAX_VALUES = SVTK_STATIC_MAX_VALUES;
constexpr long long STATIC_SCRATCH_INTS = 2 * STATIC_MAX_VALUES;

// In host code this name is a host-side stand-in, not a device address: the kernel gets the address cudaGetSymbolAddress returns.
static __device__ __align__(256) int g_static_scratch[STATIC_SCRATCH_INTS];

template <int NT>
__global__ void __launch_bounds__(NT) k_consensus_segments_static(const int32_t *__restrict__ values, const long long *__restrict__ offsets,
                                                                  const int32_t *__restrict__ pos, Tunables t, int *scratch,
                                                                  long long n_values, int cap_smem, int32_t *__restrict__ result) {
    extern __shared__ int s_dyn[];
    __shared__ ScanResult s_res[2];
    const int s = blockIdx.x;
    const long long b = offsets[s], e = offsets[s + 1];
    if (b < 0 || e < b || e > n_values) {            // a segment outside the values the caller declared: never touch the shared array
        if (threadIdx.x == 0) result[s] = SVTK_STATIC_BAD_SEGMENT;
        return;
    }
    const int n = (int)(e - b);
    if (n < t.min_count || n <= 0) { if (threadIdx.x == 0) result[s] = -1; return; }
    int *seg = (n <= cap_smem) ? s_dyn : scratch + 2 * b;
    for (int i = threadIdx.x; i < n; i += NT) seg[i] = values[b + i];
    __syncthreads();
    block_bitonic_sort(seg, n);
    const int res = block_consensus(seg, n, pos[s], t, s_res);
    if (threadIdx.x == 0) result[s] = res;
}

// the array's device address, looked up once per device with that device current
static int static_scratch(int device, int **out) {
    static std::mutex mu;
    static int *ptr[64] = {};
    if (device < 0 || device >= 64) return fail(SVTK_ERR_ARG, "device %d out of range", device);
    std::lock_guard<std::mutex> lock(mu);
    if (!ptr[device]) {
        void *p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
        if (e != cudaSuccess || !p) return fail(SVTK_ERR_CUDA, "cudaGetSymbolAddress(static scratch): %s", cudaGetErrorString(e));
        ptr[device] = (int *)p;
    }
    *out = ptr[device];
    return 0;
}

}  // namespace svtk

using namespace svtk;

extern "C" int svtk_consensus_pos_batch_static(svtk_ctx *ctx, const int32_t *dev_values, const int64_t *dev_offsets, int32_t n_seg,
                                               int64_t n_values, const int32_t *dev_pos, int32_t *dev_result) {
    Ctx *c = reinterpret_cast<Ctx *>(ctx);
    if (!c || !dev_offsets || !dev_pos || !dev_result || n_seg < 0 || n_values < 0) return fail(SVTK_ERR_ARG, "bad argument");
    if (n_values && !dev_values) return fail(SVTK_ERR_ARG, "dev_values is NULL");
    if (n_values > STATIC_MAX_VALUES)
        return fail(SVTK_ERR_ARG, "svtk_consensus_pos_batch_static: %lld values exceed the static scratch (SVTK_STATIC_MAX_VALUES = %lld); "
                                  "use svtk_consensus_pos_batch", (long long)n_values, (long long)STATIC_MAX_VALUES);
    if (n_seg == 0) return 0;
    if (int rc = set_device(c)) return rc;
    int *scratch = nullptr;
    if (int rc = static_scratch(c->device, &scratch)) return rc;
    const Tunables t{c->params.consensus_interval_range, c->params.consensus_interval, c->params.consensus_min_count};
    k_consensus_segments_static<STATIC_NT><<<n_seg, STATIC_NT, STATIC_CAP_SMEM * 4, c->stream>>>(
        dev_values, (const long long *)dev_offsets, dev_pos, t, scratch, (long long)n_values, STATIC_CAP_SMEM, dev_result);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(SVTK_ERR_CUDA, "k_consensus_segments_static: %s", cudaGetErrorString(e));
    c->launches += 1;
    return 0;
}
